// spectral front-end kernels, instantiated per frame length in lc3_k_front_*.cu (split for parallel builds)
#pragma once
#include <cuda_runtime.h>
#include "lc3_body_frame.cuh"

using namespace lc3k;

// Large blocks on purpose (here and in k_tns, k_sns, k_back): these bodies are tens of thousands of instructions of
// mostly straight-line code that every warp runs once, so the kernels are bound by instruction fetch unless the warps
// of an SM walk through the code together and share the fetches (k_sns: "no instruction" stalls 9.7 -> 0, 1.24 -> 0.89 ms).
#ifndef FRONT_WARPS
#define FRONT_WARPS 32 /* measured: 16 warps (128 registers) C3 94.6 / C4 86.4 M frames/s, 24 warps 101.2 / 93.3, 32 warps (64 registers) 104.7 / 95.6 */
#endif
#ifndef FRONT_BLOCKS
#define FRONT_BLOCKS 1
#endif
#ifndef TNS_WARPS
#define TNS_WARPS 32 /* measured at N = 480: 24 warps 1.11 ms, 32 warps (64 registers) 1.01, 12 x 2 blocks 1.12, 16 (128 registers) 1.35 */
#endif
#ifndef TNS_BLOCKS
#define TNS_BLOCKS 1
#endif
template <int N, int BITS> __global__ void __launch_bounds__(32 * FRONT_WARPS, FRONT_BLOCKS) k_front(LC3K_Batch b)
{
    const int warp = blockIdx.x * FRONT_WARPS + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (warp >= b.n_frames * b.n_tiles) return;
    front_body<N, BITS>(b, warp / b.n_tiles, warp % b.n_tiles, lane);
}

// tile path, steps 2 and 3 (lc3_body_frame.cuh): one lane per channel-frame
template <int N> __global__ void __launch_bounds__(32 * TNS_WARPS, TNS_BLOCKS) k_tns(LC3K_Batch b)
{
    const int warp = blockIdx.x * TNS_WARPS + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (warp >= b.n_frames * b.n_tiles) return;
    LC3_DBG_ENTER(0, warp);
    tns_body<N>(b, warp / b.n_tiles, warp % b.n_tiles, lane);
    LC3_DBG_LEAVE();
}

// MDCT of one tile (32 channel-frames) per block, spectrum in shared memory (lc3_mdct_tile.cuh).  In the twiddle,
// leaf, column and post phases a warp only touches its own four frames (w, w+8, w+16, w+24), so warp-level
// synchronisation is enough between them.
template <int N, int BITS> __global__ void __launch_bounds__(256, 2) k_mdct(LC3K_Batch b)
{
    typedef MdctTile<N> G;
    extern __shared__ __align__(16) i32 mt_smem[];
    unsigned long long *bar = reinterpret_cast<unsigned long long *>(mt_smem); // 16 bytes reserved in front of X
    i32 *X = mt_smem + 4, *mx = X + 32 * G::S, *mn = mx + G::PARTS * 32, *smx = mn + G::PARTS * 32;
    i32 *ben = smx + G::PARTS * 32, *bexp = ben + 32 * G::BROW;
    i16 *s_tw = (i16 *)(bexp + 32 * G::BROW), *s_sn = s_tw + G::TW_LEN, *s_W = s_sn + G::SN_LEN, *s_win = s_W + G::W_LEN;
    const int f = blockIdx.x / b.n_tiles, tile = blockIdx.x % b.n_tiles, t = threadIdx.x, w = t >> 5, l = t & 31;
    const bool rows = BITS == 16 && !b.stage_full; // PCM rows straight from the caller's buffer (bulk asynchronous copies)
    if (rows) mt_stage_rows<N>(b, f, tile, t, X, bar);
    {
        const MtTabs R = mt_rom_tabs<N>();
        for (int i = t; i < G::TW_LEN; i += G::THREADS) s_tw[i] = R.tw[i];
        for (int i = t; i < G::SN_LEN; i += G::THREADS) s_sn[i] = R.sn[i];
        for (int i = t; i < G::W_LEN; i += G::THREADS) s_W[i] = R.W[i];
        if (rows) for (int i = t; i < b.W_len; i += G::THREADS) s_win[i] = rom_win[b.W_off + i];
    }
    const MtTabs T = {s_tw, s_sn, s_W};
    LC3_DBG_ENTER(1, w == 0 ? (int)blockIdx.x : LC3_DBG_ROWS); // one row per tile, written by its first warp
    if (rows) {
        mt_wait_rows(bar);
        __syncthreads(); // the tables, too
#pragma unroll 1
        for (int j = 0; j < 4; j++) {
            const int r = w + 8 * j;
            mt_fill_row<N>(b, f, tile, r, mt_row_is_bulk<N>(b, f, tile, r), X);
            mt_fold_row<N>(b, r, tile * LC3K_LANES + r < b.n_lanes, X, mx, mn, s_win);
        }
        __syncwarp();
    } else {
        mt_fold<N, BITS>(b, f, tile, t, X, mx, mn);
        __syncthreads();
    }
#pragma unroll 1
    for (int j = 0; j < 4; j++) mt_pre<N>(X, mx, mn, w + 8 * j, l, T);
    __syncwarp();
    mt_leaf<N>(X, w + 8 * (l >> 3), l & 7);
    __syncwarp();
#pragma unroll 1
    for (int j = 0; j < 4; j++) {
        cpx o[8];
        if (l < G::LEAF) mt_col_compute<N>(X, w + 8 * j, l, o, T);
        __syncwarp();
        if (l < G::LEAF) mt_col_store<N>(X, w + 8 * j, l, o);
    }
    __syncwarp();
#pragma unroll 1
    for (int j = 0; j < 4; j++) mt_post<N>(X, w + 8 * j, l, T);
    __syncthreads();
    mt_bands<N>(b, t, X, ben, bexp, smx);
    __syncthreads();
    mt_store<N, BITS>(b, f, tile, t, X, mx, mn, ben, bexp, smx);
    LC3_DBG_LEAVE();
}

template <int N> static int front_launch(const LC3K_Batch *b, cudaStream_t s)
{
    const int warps = b->n_frames * b->n_tiles, blocks = (warps + FRONT_WARPS - 1) / FRONT_WARPS;
    if (b->bits == 16) k_front<N, 16><<<blocks, 32 * FRONT_WARPS, 0, s>>>(*b);
    else k_front<N, 24><<<blocks, 32 * FRONT_WARPS, 0, s>>>(*b);
    return (int)cudaGetLastError();
}

// frame lengths with an 8-column FFT: tile kernel (MDCT + band energies); SNS (lc3_k_sns.cu) and shaping/TNS/energies
// (tns_launch below) follow as their own launches, each one lane per channel-frame
template <int N> static int front_launch_tiled(const LC3K_Batch *b, cudaStream_t s)
{
    static bool attr_dev[64] = {false}; // the attribute is per device
    int attr_d = 0; cudaGetDevice(&attr_d); bool &attr_set = attr_dev[attr_d & 63];
    const size_t smem = MdctTile<N>::smem_bytes;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(k_mdct<N, 16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(k_mdct<N, 24>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return (int)e;
        attr_set = true;
    }
    const int tiles = b->n_frames * b->n_tiles;
    if (b->bits == 16) k_mdct<N, 16><<<tiles, 256, smem, s>>>(*b);
    else k_mdct<N, 24><<<tiles, 256, smem, s>>>(*b);
    return (int)cudaGetLastError();
}
template <int N> static int tns_launch(const LC3K_Batch *b, cudaStream_t s)
{
    const int tiles = b->n_frames * b->n_tiles, blocks = (tiles + TNS_WARPS - 1) / TNS_WARPS;
    k_tns<N><<<blocks, 32 * TNS_WARPS, 0, s>>>(*b);
    return (int)cudaGetLastError();
}
