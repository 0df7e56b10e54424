// sm_100a kernels of the batched LC3plus DECODER and their launchers (C ABI towards lc3_dec_api.c).
//
//   kd_entropy  lane per channel-frame (warp = the 32 channel-streams of a tile at one frame index): payloads staged in shared
//               memory by the warp, then side information, arithmetic decoder, residual bits and the spectrum up to SNS shaping
//   kd_idct     DCT-IV of the inverse MDCT: block per tile with the spectrum in shared memory (N >= 160), lane per channel-frame
//               (N < 160); result row-major
//   kd_synth    warp per channel-stream, frames in order: windowing + overlap-add, LTPF, output scaling, PCM to the caller
// No tensor cores: saturating 16/32-bit fixed point, as the encoder.
#include <cuda_runtime.h>
#include "lc3_dec_frame.cuh"
#include "lc3_dec_idct.cuh"
#include "lc3_dec_synth.cuh"

using namespace lc3k;

// frame bytes in shared memory, the layout of the encoder's back end (lc3_k_back.cu): word w of lane L at word w*33 + L
struct SmemBytesIn {
    const u8 *base; // warp region + 4*lane
    __device__ i32 get(int i) const { return base[(i >> 2) * 132 + (i & 3)]; }
    __device__ u32 word(int k) const { return *reinterpret_cast<const u32 *>(base + k * 132); }
};

#define DEC_WARPS 8
#define DEC_SMEM_MAX (4096 + 1088 * 4 + DEC_WARPS * (440 / 4) * 132)
static int dec_smem_bytes(int max_bytes) { return 4096 + 1088 * 4 + DEC_WARPS * ((max_bytes + 3) / 4) * 132; }
__global__ void __launch_bounds__(32 * DEC_WARPS) kd_entropy(LC3D_Batch b)
{
    extern __shared__ __align__(16) unsigned char de_smem[];
    u32 *cf = reinterpret_cast<u32 *>(de_smem);
    u8 *lut = de_smem + 1088 * 4, *frames = lut + 4096;
    for (int i = threadIdx.x; i < 4096; i += 32 * DEC_WARPS) lut[i] = (u8)rom_ari_lookup[i];
    for (int i = threadIdx.x; i < 1088; i += 32 * DEC_WARPS) cf[i] = (u32)rom_ari_freq[i] << 16 | rom_ari_cumfreq[i];
    __syncthreads();
    const int wib = threadIdx.x >> 5, warp = blockIdx.x * DEC_WARPS + wib, lane = threadIdx.x & 31;
    if (warp >= b.n_frames * b.n_tiles) return;
    const int f = warp / b.n_tiles, tile = warp % b.n_tiles;
    u8 *region = frames + (size_t)wib * ((b.max_bytes + 3) / 4) * 132;
    // copy-in by the whole warp, one payload (row) at a time
    const int li = tile * LC3K_LANES + lane;
    const u8 *my_src = nullptr;
    i32 my_n = 0;
    if (li < b.n_lanes) {
        my_n = tile_view(b.cfg, (size_t)tile, DCFG_N, lane)[DCFG_BYTES];
        my_src = b.in + b.in_off[li] + (size_t)(b.frame0 + f) * b.in_stride[li];
    }
    for (int r = 0; r < 32; r++) {
        const i32 n = __shfl_sync(0xffffffffu, my_n, r);
        const u8 *s = reinterpret_cast<const u8 *>(__shfl_sync(0xffffffffu, reinterpret_cast<unsigned long long>(my_src), r));
        if (n <= 0) continue;
        if (((reinterpret_cast<unsigned long long>(s) | (unsigned)n) & 3) == 0) {
            for (int w = lane; w < (n >> 2); w += 32) *reinterpret_cast<u32 *>(region + w * 132 + 4 * r) = reinterpret_cast<const u32 *>(s)[w];
        } else {
            for (int i = lane; i < n; i += 32) region[(i >> 2) * 132 + 4 * r + (i & 3)] = s[i];
        }
    }
    __syncwarp();
    AriTabs T = {lut, cf};
    dec_frame_body(b, f, tile, lane, T, SmemBytesIn{region + 4 * lane});
}

template <int N> __global__ void __launch_bounds__(256, 2) kd_idct_tile(LC3D_Batch b)
{
    typedef MdctTile<N> G;
    extern __shared__ __align__(16) i32 di_smem[];
    i32 *X = di_smem, *mx = X + 32 * G::S, *mn = mx + G::PARTS * 32, *red = mn + G::PARTS * 32; // red: 64 words per warp
    i16 *s_tw = (i16 *)(red + 8 * 64), *s_sn = s_tw + G::TW_LEN, *s_W = s_sn + G::SN_LEN;
    const int f = blockIdx.x / b.n_tiles, tile = blockIdx.x % b.n_tiles, t = threadIdx.x, w = t >> 5, l = t & 31;
    {
        const MtTabs R = mt_rom_tabs<N>();
        for (int i = t; i < G::TW_LEN; i += G::THREADS) s_tw[i] = R.tw[i];
        for (int i = t; i < G::SN_LEN; i += G::THREADS) s_sn[i] = R.sn[i];
        for (int i = t; i < G::W_LEN; i += G::THREADS) s_W[i] = R.W[i];
    }
    const MtTabs T = {s_tw, s_sn, s_W};
    mt_dec_load<N>(b, f, tile, t, X, mx, mn);
    __syncthreads();
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
    __syncwarp();
#pragma unroll 1
    for (int j = 0; j < 4; j++) mt_dec_store<N>(b, f, tile, w + 8 * j, X, red + 64 * w);
}
template <int N> static size_t idct_tile_smem()
{
    typedef MdctTile<N> G;
    return (size_t)(32 * G::S + 2 * G::PARTS * 32 + 8 * 64) * sizeof(i32) + (size_t)(G::TW_LEN + G::SN_LEN + G::W_LEN) * sizeof(i16);
}

// short frames: lane per channel-frame in place, then the warp transposes its 32 spectra into rows
template <int N> __global__ void __launch_bounds__(32 * DEC_WARPS) kd_idct_lane(LC3D_Batch b)
{
    __shared__ i32 tb[DEC_WARPS][32 * 33];
    const int wib = threadIdx.x >> 5, warp = blockIdx.x * DEC_WARPS + wib, lane = threadIdx.x & 31;
    if (warp >= b.n_frames * b.n_tiles) return;
    const int f = warp / b.n_tiles, tile = warp % b.n_tiles;
    idct_lane_body<N>(b, f, tile, lane);
    __syncwarp();
    const i32 *d = b.spec + dframe_tile(b, f, tile) * N * 32;
    i32 *t = tb[wib];
    for (int e0 = 0; e0 < N; e0 += 32) {
        for (int ee = 0; ee < 32; ee++) if (e0 + ee < N) t[ee * 33 + lane] = d[(size_t)(e0 + ee) * 32 + lane];
        __syncwarp();
        for (int r = 0; r < 32; r++) {
            const int li = tile * LC3K_LANES + r;
            if (li < b.n_lanes && e0 + lane < N) b.yrow[dframe_row(b, f, li) * N + e0 + lane] = t[lane * 33 + r];
        }
        __syncwarp();
    }
}

#define SYNTH_WARPS 4
struct SynthSmem { SynthShared sh[SYNTH_WARPS]; i16 win[2 * LC3K_MAX_N]; };
template <int BITS> __global__ void __launch_bounds__(32 * SYNTH_WARPS, 7) kd_synth(LC3D_Batch b)
{
    extern __shared__ __align__(16) unsigned char ds_smem[];
    SynthSmem &m = *reinterpret_cast<SynthSmem *>(ds_smem);
    for (int i = threadIdx.x; i < b.W_len; i += 32 * SYNTH_WARPS) m.win[i] = rom_win[b.W_off + i];
    __syncthreads();
    const int wib = threadIdx.x >> 5, li = blockIdx.x * SYNTH_WARPS + wib;
    if (li >= b.n_lanes) return;
    synth_warp_body<BITS>(b, li, m.sh[wib], b.yrow, m.win);
}

__global__ void kd_reset(int32_t *st32, int16_t *st16, const int32_t *slots, int n)
{
    const int i = blockIdx.x, t = threadIdx.x;
    if (i >= n) return;
    const int slot = slots[i];
    SV<i32> s = slot_view(st32, slot, DST32_N);
    SV<i16> h = slot_view(st16, slot, DST16_N);
    for (int k = t; k < DST32_N; k += blockDim.x) s[k] = k == DST_LTPF_SCALE_IDX ? -1 : 0;
    for (int k = t; k < DST16_N; k += blockDim.x) h[k] = 0;
}

static int done() { return (int)cudaGetLastError(); }
static bool *attr_flag(bool *tab) { int d = 0; cudaGetDevice(&d); return &tab[d & 63]; } // function attributes are per device

extern "C" int lc3d_launch_entropy(const LC3D_Batch *b, void *stream)
{
    static bool tab[64];
    bool *set = attr_flag(tab);
    if (!*set) {
        cudaError_t e = cudaFuncSetAttribute(kd_entropy, cudaFuncAttributeMaxDynamicSharedMemorySize, DEC_SMEM_MAX);
        if (e != cudaSuccess) return (int)e;
        *set = true;
    }
    const int warps = b->n_frames * b->n_tiles, blocks = (warps + DEC_WARPS - 1) / DEC_WARPS;
    const int smem = dec_smem_bytes(b->max_bytes);
    if (smem > DEC_SMEM_MAX) return (int)cudaErrorInvalidValue;
    kd_entropy<<<blocks, 32 * DEC_WARPS, smem, (cudaStream_t)stream>>>(*b);
    return done();
}

template <int N> static int idct_tile_launch(const LC3D_Batch *b, cudaStream_t s)
{
    static bool tab[64];
    bool *set = attr_flag(tab);
    const size_t smem = idct_tile_smem<N>();
    if (!*set) {
        cudaError_t e = cudaFuncSetAttribute(kd_idct_tile<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return (int)e;
        *set = true;
    }
    kd_idct_tile<N><<<b->n_frames * b->n_tiles, 256, smem, s>>>(*b);
    return done();
}
template <int N> static int idct_lane_launch(const LC3D_Batch *b, cudaStream_t s)
{
    const int warps = b->n_frames * b->n_tiles;
    kd_idct_lane<N><<<(warps + DEC_WARPS - 1) / DEC_WARPS, 32 * DEC_WARPS, 0, s>>>(*b);
    return done();
}
extern "C" int lc3d_launch_idct(const LC3D_Batch *b, void *stream)
{
    cudaStream_t s = (cudaStream_t)stream;
    switch (b->N) {
    case 480: return idct_tile_launch<480>(b, s);
    case 320: return idct_tile_launch<320>(b, s);
    case 240: return idct_tile_launch<240>(b, s);
    case 160: return idct_tile_launch<160>(b, s);
    case 120: return idct_lane_launch<120>(b, s);
    case 80: return idct_lane_launch<80>(b, s);
    case 60: return idct_lane_launch<60>(b, s);
    case 40: return idct_lane_launch<40>(b, s);
    case 20: return idct_lane_launch<20>(b, s);
    default: return (int)cudaErrorInvalidValue;
    }
}

extern "C" int lc3d_launch_synth(const LC3D_Batch *b, void *stream)
{
    static bool tab[64];
    bool *set = attr_flag(tab);
    if (!*set) {
        cudaError_t e = cudaFuncSetAttribute(kd_synth<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SynthSmem));
        if (e == cudaSuccess) e = cudaFuncSetAttribute(kd_synth<24>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(SynthSmem));
        if (e != cudaSuccess) return (int)e;
        *set = true;
    }
    const int blocks = (b->n_lanes + SYNTH_WARPS - 1) / SYNTH_WARPS;
    if (b->bits == 16) kd_synth<16><<<blocks, 32 * SYNTH_WARPS, sizeof(SynthSmem), (cudaStream_t)stream>>>(*b);
    else kd_synth<24><<<blocks, 32 * SYNTH_WARPS, sizeof(SynthSmem), (cudaStream_t)stream>>>(*b);
    return done();
}

extern "C" int lc3d_launch_reset(int32_t *st32, int16_t *st16, const int32_t *slots, int n, void *stream)
{
    if (n <= 0) return 0;
    kd_reset<<<n, 128, 0, (cudaStream_t)stream>>>(st32, st16, slots, n);
    return done();
}
