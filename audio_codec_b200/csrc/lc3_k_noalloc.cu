// lc3plus_enc_fresh_noalloc: the encode chain for callers that may not allocate (CUDA graph capture, tight serving
// loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
//
// Every byte the chain's kernels use as scratch -- the per-stream state (power-on values), the lane tables and the nine
// per-chunk work buffers that lc3_api.c otherwise cudaMallocs -- is carved out of ONE file-scope device array of fixed
// size.  The entry point only queues kernels and one memset on the caller's stream: no cudaMalloc / cudaFree / pinned
// allocation, no host<->device copy, no synchronisation.  The array is shared by every call (per GPU) and keeps the last
// call's contents: state is cleared at the start of each call, and two calls that overlap on the GPU race.
#include <cuda_runtime.h>
#include <stdio.h>
#include <string.h>
#include "../../include/lc3plus_b200.h"
extern "C" {
#include "lc3_config.h"
void lc3b_set_error(const char *msg); // lc3_api.c: text returned by lc3plus_last_error
}

#define NOALLOC_SCRATCH_BYTES (7u << 20) /* 7 MiB: LC3PLUS_NOALLOC_MAX_CHANNEL_FRAMES frames of 480 samples at 24 bit */
static __device__ __align__(256) unsigned char g_noalloc_scratch[NOALLOC_SCRATCH_BYTES];

// The array's name is a host-side stand-in in host code: its device address is looked up once per device, with that
// device current, and that pointer is what the kernels get.
#define NOALLOC_MAX_DEV 64
static unsigned char *g_scratch_dev[NOALLOC_MAX_DEV];

struct NoallocCfg { int32_t w[CFG_N]; };

// lane tables, configuration words and the power-on constants of the state (everything else in the state region was
// zeroed by the memset queued before this kernel): one thread per slot
__global__ void k_noalloc_setup(int n_lanes, int n_slots, int32_t *slot, int32_t *ostride, int64_t *poff, int64_t *ooff,
                                int32_t *cfg, int32_t *st32, NoallocCfg c, int64_t pcm_stride_elems, int64_t out_stride, int nbytes)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_slots) return;
    const size_t tile = (size_t)(i >> 5), lane = (size_t)(i & 31);
    for (int k = 0; k < CFG_N; k++) cfg[(tile * CFG_N + k) * 32 + lane] = c.w[k];
    st32[(tile * ST32_N + ST_X_EXP) * 32 + lane] = 15;      // fx/setup_enc_lc3.c:38-66
    st32[(tile * ST32_N + ST_RESAMP_EXP) * 32 + lane] = 17;
    if (i < n_lanes) {
        slot[i] = i; ostride[i] = nbytes;
        poff[i] = (int64_t)i * pcm_stride_elems;
        ooff[i] = (int64_t)i * out_stride;
    }
}

static LC3_Error fail(const char *msg) { lc3b_set_error(msg); return LC3_ERROR; }

static int is_dev_ptr(const void *p)
{
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return 0; }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

extern "C" LC3_Error lc3plus_enc_fresh_noalloc(int samplerate, int frame_dms, int bitrate, int bitdepth, int n_streams,
                                               int n_frames, const void *d_pcm, size_t stream_stride_bytes, void *d_out,
                                               size_t out_stream_stride_bytes, void *cuda_stream)
{
    if (!d_pcm || !d_out) return LC3_NULL_ERROR;
    if (bitdepth != 16 && bitdepth != 24) return fail("lc3plus_enc_fresh_noalloc: bitdepth must be 16 or 24");
    LC3Cfg c;
    int r = lc3cfg_init(&c, samplerate, 1);
    if (!r) r = lc3cfg_set_frame_dms(&c, frame_dms);
    if (!r) r = lc3cfg_set_bitrate(&c, bitrate);
    if (r) return (LC3_Error)r;
    if (n_streams <= 0 || n_frames <= 0) return fail("lc3plus_enc_fresh_noalloc: n_streams and n_frames must be positive");
    const int n_tiles = (n_streams + 31) / 32, n_slots = n_tiles * 32;
    if (n_streams > LC3PLUS_NOALLOC_MAX_STREAMS || (long long)n_slots * n_frames > LC3PLUS_NOALLOC_MAX_CHANNEL_FRAMES) {
        char m[200];
        snprintf(m, sizeof(m), "lc3plus_enc_fresh_noalloc: %d streams x %d frames is beyond the static scratch (at most %d streams and "
                 "%d channel-frames, streams rounded up to a multiple of 32)", n_streams, n_frames, LC3PLUS_NOALLOC_MAX_STREAMS,
                 LC3PLUS_NOALLOC_MAX_CHANNEL_FRAMES);
        return fail(m);
    }
    const int N = c.N, sb = bitdepth == 16 ? 2 : 4, nbytes = lc3cfg_num_bytes(&c);
    if (stream_stride_bytes % sb || (n_streams > 1 && (stream_stride_bytes < (size_t)n_frames * N * sb || out_stream_stride_bytes < (size_t)n_frames * nbytes)))
        return fail("lc3plus_enc_fresh_noalloc: stream stride too small or misaligned");
    if (!is_dev_ptr(d_pcm) || !is_dev_ptr(d_out)) return fail("lc3plus_enc_fresh_noalloc: pcm and out must be device pointers");

    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= NOALLOC_MAX_DEV) return fail("lc3plus_enc_fresh_noalloc: no CUDA device");
    if (!g_scratch_dev[dev]) { // once per device, with that device current
        void *p = NULL;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) {
            char m[200];
            snprintf(m, sizeof(m), "lc3plus_enc_fresh_noalloc: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
            return fail(m);
        }
        g_scratch_dev[dev] = (unsigned char *)p;
    }
    unsigned char *base = g_scratch_dev[dev];
    size_t off = 0;
    auto carve = [&](size_t bytes) { unsigned char *p = base + off; off += (bytes + 255) & ~(size_t)255; return p; };

    LC3K_Batch b;
    memset(&b, 0, sizeof(b));
    lc3cfg_fill_batch(&c, &b);
    b.stage_full = !(bitdepth == 16 && c.N >= 160);
    b.bits = bitdepth; b.n_lanes = n_streams; b.n_tiles = n_tiles; b.n_frames = n_frames; b.frame0 = 0;
    b.max_bytes = c.ch[0][CFG_TARGET_BYTES];
    // state region first (zeroed as one block)
    int32_t *st32 = (int32_t *)carve((size_t)n_slots * ST32_N * 4);
    int16_t *st16 = (int16_t *)carve((size_t)n_slots * ST16_N * 2);
    int16_t *hist16 = (int16_t *)carve((size_t)n_slots * LC3K_MAX_N * 2);
    int32_t *hist32 = (int32_t *)carve((size_t)n_slots * LC3K_MAX_N * 4);
    const size_t state_bytes = off;
    int32_t *cfgw = (int32_t *)carve((size_t)n_slots * CFG_N * 4);
    int32_t *slot = (int32_t *)carve((size_t)n_slots * 4), *ostride = (int32_t *)carve((size_t)n_slots * 4);
    int64_t *poff = (int64_t *)carve((size_t)n_slots * 8), *ooff = (int64_t *)carve((size_t)n_slots * 8);
    const size_t ftl = (size_t)n_frames * n_slots;
    b.pcm_t = carve((size_t)n_slots * (n_frames + 1) * N * sb);
    b.pout = (int32_t *)carve(ftl * PO_N * 4);
    b.spec = (int32_t *)carve(ftl * N * 4);
    b.fout = (int32_t *)carve(ftl * FO_N * 4);
    b.en4 = (int32_t *)carve(ftl * (N / 4) * 4);
    b.rout = (int32_t *)carve(ftl * RO_N * 4);
    b.r12 = (int16_t *)carve(ftl * 128 * 2);
    b.xq = (int16_t *)carve(ftl * N * 2);
    b.ben = (int32_t *)carve(ftl * BEN_N * 4);
    if (off > NOALLOC_SCRATCH_BYTES) return fail("lc3plus_enc_fresh_noalloc: internal: static scratch too small for the stated limits");
    b.slot = slot; b.cfg = cfgw; b.st32 = st32; b.st16 = st16; b.hist16 = hist16; b.hist32 = hist32;
    b.pcm = d_pcm; b.pcm_off = poff; b.out = (uint8_t *)d_out; b.out_off = ooff; b.out_stride = ostride; b.bfi = NULL;

    cudaStream_t st = (cudaStream_t)cuda_stream;
    if (cudaMemsetAsync(base, 0, state_bytes, st) != cudaSuccess) return fail("lc3plus_enc_fresh_noalloc: cudaMemsetAsync failed");
    NoallocCfg cw;
    for (int k = 0; k < CFG_N; k++) cw.w[k] = c.ch[0][k];
    k_noalloc_setup<<<(n_slots + 63) / 64, 64, 0, st>>>(n_streams, n_slots, slot, ostride, poff, ooff, cfgw, st32, cw,
                                                         (int64_t)(stream_stride_bytes / sb), (int64_t)out_stream_stride_bytes, nbytes);
    if (cudaGetLastError() != cudaSuccess) return fail("lc3plus_enc_fresh_noalloc: setup launch failed");
    int (*const chain[10])(const LC3K_Batch *, void *) = {lc3k_launch_stage, lc3k_launch_fir, lc3k_launch_scalar, lc3k_launch_pitch, lc3k_launch_front,
                                                          lc3k_launch_sns, lc3k_launch_tns, lc3k_launch_rate, lc3k_launch_quant, lc3k_launch_back};
    for (int k = 0; k < 10; k++) {
        const int e = chain[k](&b, st);
        if (e) {
            char m[200];
            snprintf(m, sizeof(m), "lc3plus_enc_fresh_noalloc: kernel %d: %s", k, cudaGetErrorString((cudaError_t)e));
            return fail(m);
        }
    }
    return LC3_OK;
}
