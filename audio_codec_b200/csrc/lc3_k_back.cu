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

// Frame bytes in shared memory: word w of lane L sits at word w*33 + L of the warp's region.  A lane's bytes are always
// in "its" bank column shifted by w, so accesses with one index for the whole warp never conflict, lane-varying indices
// (the two writers of different frames sit at different bytes) conflict only by chance, and a whole row can be read by the
// 32 lanes (w = lane) without conflict for the copy-out.
struct SmemBytes {
    u8 *base; // warp region + 4*lane
    __device__ i32 get(int i) const { return base[(i >> 2) * 132 + (i & 3)]; }
    __device__ void set(int i, i32 v) const { base[(i >> 2) * 132 + (i & 3)] = (u8)v; }
    __device__ void zero(int n) const { for (int w = 0; w < (n + 3) >> 2; w++) *reinterpret_cast<u32 *>(base + w * 132) = 0; }
};

// small blocks here, unlike k_front / k_sns / k_tns: the warps of this kernel run for very different times (the entropy
// coder loops are data dependent), and a large block holds its SM slot until its slowest warp is done
#ifndef BACK_WARPS
#define BACK_WARPS 8
#endif
#define BACK_SMEM_MAX (4096 + 1088 * 4 + BACK_WARPS * (440 / 4) * 132) /* frames of up to 440 bytes */
static int back_smem_bytes(int max_bytes) { return 4096 + 1088 * 4 + BACK_WARPS * ((max_bytes + 3) / 4) * 132; }
__global__ void __launch_bounds__(32 * BACK_WARPS) k_back(LC3K_Batch b)
{
    extern __shared__ __align__(16) unsigned char bk_smem[];
    u32 *cf = reinterpret_cast<u32 *>(bk_smem);
    u8 *lut = bk_smem + 1088 * 4, *frames = lut + 4096;
    for (int i = threadIdx.x; i < 4096; i += 32 * BACK_WARPS) lut[i] = (u8)rom_ari_lookup[i];
    for (int i = threadIdx.x; i < 1088; i += 32 * BACK_WARPS) cf[i] = (u32)rom_ari_freq[i] << 16 | rom_ari_cumfreq[i];
    __syncthreads();
    const int wib = threadIdx.x >> 5, warp = blockIdx.x * BACK_WARPS + wib, lane = threadIdx.x & 31;
    if (warp >= b.n_frames * b.n_tiles) return;
    AriTabs T = {lut, cf};
    LC3_DBG_ENTER(1, warp);
    u8 *region = frames + (size_t)wib * ((b.max_bytes + 3) / 4) * 132;
    u8 *dst;
    const i32 nb = back_body(b, warp / b.n_tiles, warp % b.n_tiles, lane, T, SmemBytes{region + 4 * lane}, dst);
    __syncwarp();
    // copy-out by the whole warp, one frame (row) at a time
    for (int r = 0; r < 32; r++) {
        const i32 n = __shfl_sync(0xffffffffu, nb, r);
        u8 *d = reinterpret_cast<u8 *>(__shfl_sync(0xffffffffu, reinterpret_cast<unsigned long long>(dst), r));
        if (n == 0) continue;
        if (((reinterpret_cast<unsigned long long>(d) | (unsigned)n) & 3) == 0) {
            for (int w = lane; w < (n >> 2); w += 32) reinterpret_cast<u32 *>(d)[w] = *reinterpret_cast<const u32 *>(region + w * 132 + 4 * r);
        } else {
            for (int i = lane; i < n; i += 32) d[i] = region[(i >> 2) * 132 + 4 * r + (i & 3)];
        }
    }
    LC3_DBG_LEAVE();
}
LC3_DBG_READER(back)

// error protection of the finished payloads (lc3_fec.cuh), lane per channel-frame, in place in the caller's output
__global__ void __launch_bounds__(128) k_fec(LC3K_Batch b)
{
    __shared__ FecTabs T; // GF(16) products and CRC masks, built once per block
    for (int i = threadIdx.x; i < 256 + 48; i += 128) fec_tabs_fill(T, i);
    __syncthreads();
    const int warp = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (warp >= b.n_frames * b.n_tiles) return;
    fec_body(b, warp / b.n_tiles, warp % b.n_tiles, lane, T);
}

static int done() { return (int)cudaGetLastError(); }

extern "C" int lc3k_launch_quant(const LC3K_Batch *b, void *stream)
{
    static bool attr_dev[64] = {false}; // the attribute is per device
    int attr_d = 0; cudaGetDevice(&attr_d); bool &attr_set = attr_dev[attr_d & 63];
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(k_quant, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(QuantSmem));
        if (e != cudaSuccess) return (int)e;
        attr_set = true;
    }
    k_quant<<<b->n_frames * b->n_tiles, QT_THREADS, sizeof(QuantSmem), (cudaStream_t)stream>>>(*b);
    return done();
}
extern "C" int lc3k_launch_fec(const LC3K_Batch *b, void *stream)
{
    const int warps = b->n_frames * b->n_tiles;
    k_fec<<<(warps + 3) / 4, 128, 0, (cudaStream_t)stream>>>(*b);
    return done();
}
extern "C" int lc3k_launch_back(const LC3K_Batch *b, void *stream)
{
    static bool attr_dev[64] = {false}; // the attribute is per device
    int attr_d = 0; cudaGetDevice(&attr_d); bool &attr_set = attr_dev[attr_d & 63];
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(k_back, cudaFuncAttributeMaxDynamicSharedMemorySize, BACK_SMEM_MAX);
        if (e != cudaSuccess) return (int)e;
        attr_set = true;
    }
    const int warps = b->n_frames * b->n_tiles, blocks = (warps + BACK_WARPS - 1) / BACK_WARPS;
    const int smem = back_smem_bytes(b->max_bytes);
    if (smem > BACK_SMEM_MAX) return (int)cudaErrorInvalidValue;
    k_back<<<blocks, 32 * BACK_WARPS, smem, (cudaStream_t)stream>>>(*b);
    return done();
}
