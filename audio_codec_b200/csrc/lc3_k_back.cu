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
#include "lc3_quant_tile.cuh"

using namespace lc3k;

// gain adjustment + final quantisation of one tile (32 channel-frames) per block (lc3_quant_tile.cuh)
struct QuantSmem { QuantShared sh; uint16_t bits[1088]; u8 lut[4096]; }; // > 48 kB: dynamic shared memory
__global__ void __launch_bounds__(QT_THREADS) k_quant(LC3K_Batch b)
{
    extern __shared__ __align__(16) unsigned char qt_smem[];
    QuantSmem &m = *reinterpret_cast<QuantSmem *>(qt_smem);
    QuantShared &sh = m.sh;
    const int t = threadIdx.x, w = t >> 5;
    for (int i = t; i < 4096; i += QT_THREADS) m.lut[i] = (u8)rom_ari_lookup[i];
    for (int i = t; i < 1088; i += QT_THREADS) m.bits[i] = rom_ari_bits[i];
    const int f = blockIdx.x / b.n_tiles, tile = blockIdx.x % b.n_tiles;
    LC3_DBG_ENTER(0, w == 0 ? (int)blockIdx.x : LC3_DBG_ROWS);
    if (t < 32) qt_setup(b, f, tile, t, sh.fr[t]);
    __syncthreads();
    qt_quantise(b, f, tile, t, sh);
    __syncthreads();
#pragma unroll 1
    for (int j = 0; j < 4; j++) qt_analyse(sh.fr[w + 8 * j], sh.xq + (w + 8 * j) * QT_ROW, b.yLen, sh.w[w], m.lut, m.bits);
    __syncthreads();
    qt_store(b, f, tile, t, sh);
    LC3_DBG_LEAVE();
}

// small blocks here, unlike k_front / k_sns / k_tns: the warps of this kernel run for very different times (the entropy
// coder loops are data dependent), and a large block holds its SM slot until its slowest warp is done (measured: 2.44 ms
// at 4 warps per block, 2.76 at 12, 3.52 at 24)
#define BACK_WARPS 4
__global__ void __launch_bounds__(32 * BACK_WARPS, 12) k_back(LC3K_Batch b)
{
    __shared__ u8 lut[4096];
    __shared__ uint16_t cum[1088], freq[1088];
    for (int i = threadIdx.x; i < 4096; i += 32 * BACK_WARPS) lut[i] = (u8)rom_ari_lookup[i];
    for (int i = threadIdx.x; i < 1088; i += 32 * BACK_WARPS) { cum[i] = rom_ari_cumfreq[i]; freq[i] = rom_ari_freq[i]; }
    __syncthreads();
    const int warp = blockIdx.x * BACK_WARPS + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (warp >= b.n_frames * b.n_tiles) return;
    AriTabs T = {lut, cum, freq};
    LC3_DBG_ENTER(1, warp);
    back_body(b, warp / b.n_tiles, warp % b.n_tiles, lane, T);
    LC3_DBG_LEAVE();
}
LC3_DBG_READER(back)

static int done() { return (int)cudaGetLastError(); }

extern "C" int lc3k_launch_quant(const LC3K_Batch *b, void *stream)
{
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(k_quant, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(QuantSmem));
        if (e != cudaSuccess) return (int)e;
        attr_set = true;
    }
    k_quant<<<b->n_frames * b->n_tiles, QT_THREADS, sizeof(QuantSmem), (cudaStream_t)stream>>>(*b);
    return done();
}
extern "C" int lc3k_launch_back(const LC3K_Batch *b, void *stream)
{
    const int warps = b->n_frames * b->n_tiles, blocks = (warps + BACK_WARPS - 1) / BACK_WARPS;
    k_back<<<blocks, 32 * BACK_WARPS, 0, (cudaStream_t)stream>>>(*b);
    return done();
}
