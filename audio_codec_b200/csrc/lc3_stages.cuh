// Kernel bodies of the batched encoder.  Each body is the work of ONE lane; the __global__ wrappers in
// lc3_kernels.cu map threads to (tile, lane[, frame]) and call them.  (tests/hostsim calls the same
// bodies from plain loops to debug without a GPU.)
//
// Pipeline per chunk of frames (stage order of fx/enc_lc3.c:64-261, regrouped by dependency):
//   pitch_body  lane = channel-stream, walks frames in order: [24-bit block scaling] resampler, high-pass,
//               OLPA, LTPF, attack detector                                   (inter-frame state)
//   front_body  lane = channel-frame: MDCT, band energy, bandwidth detector, SNS, shaping, TNS,
//               per-quadruple log-energies for the gain estimate             (frame-parallel)
//   rate_body   lane = channel-stream, walks frames in order: gain bisection + first-pass bit count
//               (targetBitsOff / mem_targetBits / mem_specBits feedback)      (inter-frame state)
//   back_body   lane = channel-frame: gain adjust, final quantisation, residual, noise factor, side info,
//               arithmetic coder                                              (frame-parallel)
#pragma once
#include <type_traits>
#include "lc3_layout.h"
#include "lc3_mdct.cuh"
#include "lc3_pitch.cuh"
#include "lc3_spec.cuh"
#include "lc3_sns.cuh"
#include "lc3_tns.cuh"
#include "lc3_rate.cuh"
#include "lc3_bits.cuh"

namespace lc3k {

// strided view of one lane's array inside a tile-interleaved buffer
template <class T> struct SV {
    T *p;
    LC3_HDM T &operator[](int i) const { return p[(size_t)i * LC3K_LANES]; }
    LC3_HDM SV operator+(int i) const { SV r = {p + (size_t)i * LC3K_LANES}; return r; }
};
template <class T> LC3_HD SV<T> tile_view(T *base, size_t tile, size_t n_elems, int lane)
{
    SV<T> r = {base + tile * n_elems * LC3K_LANES + lane};
    return r;
}
// state pool addressed by slot
template <class T> LC3_HD SV<T> slot_view(T *base, int slot, size_t n_elems)
{
    return tile_view(base, (size_t)(slot >> 5), n_elems, slot & 31);
}

// scaled 16-bit sample access into the staged PCM of one lane: index i relative to the start of frame f,
// i in [-N, N).  16-bit input is used as is; 24-bit input is block-scaled by `sh` (scale_signal24_fx.c:60-75).
template <int BITS> struct PcmView;
template <> struct PcmView<16> {
    SV<const i16> p; // points at sample 0 of the frame
    i32 sh;
    LC3_HDM i32 operator()(int i) const { return p[i]; }
};
template <> struct PcmView<24> {
    SV<const i32> p;
    i32 sh;
    LC3_HDM i32 operator()(int i) const { return l_round(l_shl(p[i], sh)); }
};
template <int BITS> LC3_HD PcmView<BITS> pcm_view(const LC3K_Batch &b, int tile, int lane, int f, i32 x_exp)
{
    typedef typename std::conditional<BITS == 16, const i16, const i32>::type S;
    const size_t per_tile = (size_t)(b.n_frames + 1) * b.N;
    PcmView<BITS> v;
    v.p = tile_view((S *)b.pcm_t, (size_t)tile, per_tile, lane) + (f + 1) * b.N;
    v.sh = s_sub(23, x_exp);
    return v;
}

LC3_HD size_t frame_tile(const LC3K_Batch &b, int f, int tile) { return (size_t)f * b.n_tiles + tile; }

} // namespace lc3k
