// Shared by the decoder's passes (lc3_dec_frame.cuh, lc3_dec_idct.cuh, lc3_dec_synth.cuh).
#pragma once
#include "lc3_dec_layout.h"
#include "lc3_stages.cuh"
#include "lc3_warp.cuh"
#include "lc3_dec_rom.cuh"

namespace lc3k {
LC3_HD size_t dframe_tile(const LC3D_Batch &b, int f, int tile) { return (size_t)f * b.n_tiles + tile; }
// row of channel-frame (f, batch lane li) in the row-major buffer of DCT-IV outputs
LC3_HD size_t dframe_row(const LC3D_Batch &b, int f, int li) { return (size_t)f * b.n_tiles * 32 + li; }
} // namespace lc3k
