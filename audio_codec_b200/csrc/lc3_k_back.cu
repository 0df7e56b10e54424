// sm_100a kernels of the batched LC3plus encoder and their launchers (C ABI towards lc3_api.c), part: back end (quantisation, residual, noise factor, entropy coding).
//
// Thread mapping: frame-parallel passes run one lane per channel-frame (32 lanes of a warp = the same 32
// channel-streams of a tile at one frame index); per-stream passes run one warp per channel-stream.  Every
// per-lane array access with a warp-uniform index is one coalesced request into the tile-interleaved
// buffers (lc3_layout.h).  No tensor cores: the arithmetic is saturating 16/32-bit fixed point.
#include <cuda_runtime.h>
#include "lc3_body_frame.cuh"
#include "lc3_pitch_warp.cuh"
#include "lc3_rate_warp.cuh"

using namespace lc3k;

__global__ void __launch_bounds__(128) k_back(LC3K_Batch b)
{
    __shared__ u8 lut[4096];
    __shared__ uint16_t cum[1088], freq[1088], bits[1088];
    for (int i = threadIdx.x; i < 4096; i += 128) lut[i] = (u8)rom_ari_lookup[i];
    for (int i = threadIdx.x; i < 1088; i += 128) { cum[i] = rom_ari_cumfreq[i]; freq[i] = rom_ari_freq[i]; bits[i] = rom_ari_bits[i]; }
    __syncthreads();
    const int warp = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (warp >= b.n_frames * b.n_tiles) return;
    AriTabs T = {lut, cum, freq};
    back_body(b, warp / b.n_tiles, warp % b.n_tiles, lane, T, bits);
}

static int done() { return (int)cudaGetLastError(); }

extern "C" int lc3k_launch_back(const LC3K_Batch *b, void *stream)
{
    const int warps = b->n_frames * b->n_tiles, blocks = (warps + 3) / 4;
    k_back<<<blocks, 128, 0, (cudaStream_t)stream>>>(*b);
    return done();
}
