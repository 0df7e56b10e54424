// sm_100a kernels of the batched LC3plus encoder and their launchers (C ABI towards lc3_api.c), part: staging, per-stream time-domain pass, per-stream rate-loop pass, state reset.
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

// ---- staging: caller's PCM (sample-major per stream) -> tile-interleaved, history frame first.
// grid = (tiles, segments of STAGE_SEG samples); 32 lanes x 32 samples per step through shared memory.
#define STAGE_SEG 1024
template <class T> __global__ void __launch_bounds__(256) k_stage(LC3K_Batch b)
{
    __shared__ T sm[32][33];
    const int tile = blockIdx.x, tx = threadIdx.x & 31, ty = threadIdx.x >> 5; // 32 x 8
    const int N = b.N, total = b.n_frames * N;
    const size_t per_tile = (size_t)(b.n_frames + 1) * N;
    T *dst = (T *)b.pcm_t + (size_t)tile * per_tile * LC3K_LANES;
    const T *src = (const T *)b.pcm + (size_t)b.frame0 * N;

    if (blockIdx.y == 0) { // history prefix from the state, then the new history = last frame of the chunk (lane = tx)
        T *hist_pool = sizeof(T) == 2 ? (T *)b.hist16 : (T *)b.hist32;
        const int li = tile * 32 + tx;
        if (li < b.n_lanes) {
            SV<T> h = slot_view(hist_pool, b.slot[li], LC3K_MAX_N);
            const T *last = src + b.pcm_off[li] + (total - N);
            for (int i = ty; i < N; i += 8) { dst[(size_t)i * 32 + tx] = h[i]; h[i] = last[i]; }
        }
    }
    if (!b.stage_full) return; // history only: every later frame is read from the caller's PCM where it lies
    const int c_end = min(total, (int)(blockIdx.y + 1) * STAGE_SEG);
    for (int c0 = blockIdx.y * STAGE_SEG; c0 < c_end; c0 += 32) {
#pragma unroll
        for (int r = 0; r < 4; r++) {
            const int lane = ty + 8 * r, li = tile * 32 + lane, i = c0 + tx;
            T v = 0;
            if (li < b.n_lanes && i < total) v = src[b.pcm_off[li] + i];
            sm[lane][tx] = v;
        }
        __syncthreads();
#pragma unroll
        for (int r = 0; r < 4; r++) {
            const int k = ty + 8 * r, i = c0 + k;
            if (i < total) dst[(size_t)(N + i) * 32 + tx] = sm[tx][k];
        }
        __syncthreads();
    }
}

// per-stream passes: one warp per channel-stream.  14 warps per block, 2 blocks per SM: 28 x 148 = 4144 streams are
// resident at once (the 4096-stream workload runs in one wave), and the warps of a block share instruction fetches.
#ifndef PITCH_WARPS
#define PITCH_WARPS 14
#endif
#ifndef PITCH_FRAME_SYNC
#define PITCH_FRAME_SYNC 1
#endif
template <int BITS, bool PRE, int DMS> __global__ void __launch_bounds__(32 * PITCH_WARPS, 28 / PITCH_WARPS) k_pitch(LC3K_Batch b)
{
    extern __shared__ __align__(16) unsigned char pitch_smem[];
    PitchShared *sh = reinterpret_cast<PitchShared *>(pitch_smem);
    const int w = threadIdx.x >> 5;
    LC3_DBG_ENTER(0, (int)blockIdx.x * PITCH_WARPS + w);
    const int live = min(PITCH_WARPS, b.n_lanes - (int)blockIdx.x * PITCH_WARPS); // warps of this block that own a stream
    pitch_warp_body<BITS, PRE, DMS>(b, blockIdx.x * PITCH_WARPS + w, sh[w], PITCH_FRAME_SYNC ? 32 * live : 0);
    LC3_DBG_LEAVE();
}

// resampler FIR, frame-parallel: one warp per (channel-stream, frame)
// A block's warps each take FIR_ITEMS (stream, frame) pairs, so the tap tables (read lane by lane from constant
// memory, which serialises) are set up once per FIR_WARPS * FIR_ITEMS frames.
#define FIR_WARPS 8
#define FIR_ITEMS 4
__global__ void __launch_bounds__(32 * FIR_WARPS) k_fir(LC3K_Batch b)
{
    __shared__ FirShared sh[FIR_WARPS];
    __shared__ i16 filt[240];
    __shared__ u32 filtp[120];
    for (int i = threadIdx.x; i < 240; i += 32 * FIR_WARPS) filt[i] = rom_resamp_filt[b.fs_idx * 240 + i];
    __syncthreads();
    if (threadIdx.x < 120) filtp[threadIdx.x] = fir_pack_taps(filt[2 * threadIdx.x], filt[2 * threadIdx.x + 1]);
    __syncthreads();
    const int w = threadIdx.x >> 5;
    const long long n_items = (long long)b.n_frames * b.n_lanes;
    for (int k = 0; k < FIR_ITEMS; k++) {
        const long long item = ((long long)blockIdx.x * FIR_ITEMS + k) * FIR_WARPS + w;
        if (item >= n_items) return;
        fir_warp_body(b, (int)(item % b.n_lanes), (int)(item / b.n_lanes), sh[w], filt, filtp);
        __syncwarp();
    }
}

// scalar recursions (high-pass, attack detector): one lane per channel-stream
__global__ void __launch_bounds__(32) k_scalar(LC3K_Batch b) { scalar_body(b, blockIdx.x, threadIdx.x); }

// rate loop: one warp per channel-stream, like k_pitch; the block's warps meet before every frame (shared instruction fetches)
#ifndef RATE_WARPS
#define RATE_WARPS 14
#endif
__global__ void __launch_bounds__(32 * RATE_WARPS, 896 / (32 * RATE_WARPS)) k_rate(LC3K_Batch b)
{
    __shared__ RateShared sh[RATE_WARPS];
    __shared__ u8 lut[4096];
    __shared__ uint16_t bits[1088];
    for (int i = threadIdx.x; i < 4096; i += 32 * RATE_WARPS) lut[i] = (u8)rom_ari_lookup[i];
    for (int i = threadIdx.x; i < 1088; i += 32 * RATE_WARPS) bits[i] = rom_ari_bits[i];
    __syncthreads();
    const int w = threadIdx.x >> 5;
    LC3_DBG_ENTER(1, (int)blockIdx.x * RATE_WARPS + w);
    const int live = min(RATE_WARPS, b.n_lanes - (int)blockIdx.x * RATE_WARPS);
    rate_warp_body(b, blockIdx.x * RATE_WARPS + w, sh[w], lut, bits, 32 * live);
    LC3_DBG_LEAVE();
}
LC3_DBG_READER(stream)

__global__ void k_reset(i32 *st32, i16 *st16, i16 *hist16, i32 *hist32, const i32 *slots, int n)
{
    const int k = blockIdx.x;
    if (k >= n) return;
    const int slot = slots[k];
    SV<i32> a = slot_view(st32, slot, ST32_N);
    SV<i16> c = slot_view(st16, slot, ST16_N);
    SV<i16> h = slot_view(hist16, slot, LC3K_MAX_N);
    SV<i32> g = slot_view(hist32, slot, LC3K_MAX_N);
    for (int i = threadIdx.x; i < ST32_N; i += blockDim.x) a[i] = i == ST_X_EXP ? 15 : (i == ST_RESAMP_EXP ? 17 : 0);
    for (int i = threadIdx.x; i < ST16_N; i += blockDim.x) c[i] = 0;
    for (int i = threadIdx.x; i < LC3K_MAX_N; i += blockDim.x) { h[i] = 0; g[i] = 0; }
}

static int done() { return (int)cudaGetLastError(); }

extern "C" int lc3k_launch_reset(int32_t *st32, int16_t *st16, int16_t *hist16, int32_t *hist32, const int32_t *slots, int n, void *stream)
{
    if (n <= 0) return 0;
    k_reset<<<n, 64, 0, (cudaStream_t)stream>>>(st32, st16, hist16, hist32, slots, n);
    return done();
}
extern "C" int lc3k_launch_stage(const LC3K_Batch *b, void *stream)
{
    const dim3 grid(b->n_tiles, b->stage_full ? (b->n_frames * b->N + STAGE_SEG - 1) / STAGE_SEG : 1);
    if (b->bits == 16) k_stage<int16_t><<<grid, 256, 0, (cudaStream_t)stream>>>(*b);
    else k_stage<int32_t><<<grid, 256, 0, (cudaStream_t)stream>>>(*b);
    return done();
}
// 16-bit input: FIR (frame-parallel), high-pass recursion (lane per stream), then pitch analysis (warp per stream);
// 24-bit input: the fused kernel does all three (lc3k_launch_fir / _scalar are then no-ops)
extern "C" int lc3k_launch_fir(const LC3K_Batch *b, void *stream)
{
    if (b->bits != 16) return 0;
    const long long items = (long long)b->n_frames * b->n_lanes;
    k_fir<<<(unsigned)((items + FIR_WARPS * FIR_ITEMS - 1) / (FIR_WARPS * FIR_ITEMS)), 32 * FIR_WARPS, 0, (cudaStream_t)stream>>>(*b);
    return done();
}
extern "C" int lc3k_launch_scalar(const LC3K_Batch *b, void *stream)
{
    if (b->bits != 16) return 0;
    k_scalar<<<b->n_tiles, 32, 0, (cudaStream_t)stream>>>(*b);
    return done();
}
template <int BITS, bool PRE, int DMS> static int pitch_launch(const LC3K_Batch *b, cudaStream_t stream)
{
    static bool attr_dev[64] = {false}; // the attribute is per device
    int attr_d = 0; cudaGetDevice(&attr_d); bool &attr_set = attr_dev[attr_d & 63];
    const size_t smem = PITCH_WARPS * sizeof(PitchShared);
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(k_pitch<BITS, PRE, DMS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return (int)e;
        attr_set = true;
    }
    k_pitch<BITS, PRE, DMS><<<(b->n_lanes + PITCH_WARPS - 1) / PITCH_WARPS, 32 * PITCH_WARPS, smem, stream>>>(*b);
    return done();
}
extern "C" int lc3k_launch_pitch(const LC3K_Batch *b, void *stream)
{ // one instance per input width and frame duration
    cudaStream_t s = (cudaStream_t)stream;
    if (b->bits == 16) return b->dms == 100 ? pitch_launch<16, true, 100>(b, s) : b->dms == 50 ? pitch_launch<16, true, 50>(b, s) : pitch_launch<16, true, 25>(b, s);
    return b->dms == 100 ? pitch_launch<24, false, 100>(b, s) : b->dms == 50 ? pitch_launch<24, false, 50>(b, s) : pitch_launch<24, false, 25>(b, s);
}
extern "C" int lc3k_launch_rate(const LC3K_Batch *b, void *stream)
{
    k_rate<<<(b->n_lanes + RATE_WARPS - 1) / RATE_WARPS, 32 * RATE_WARPS, 0, (cudaStream_t)stream>>>(*b);
    return done();
}
