/* Host C layer of liblc3plus_b200: the reference encoder API (fx/lc3.h, fx/lc3.c:94-267) and the batched
 * entry point on top of the CUDA kernels in lc3_kernels.cu.  Plain C; talks to the device through the
 * CUDA runtime C API and the lc3k_launch_* functions.  There is no CPU encode path in this library.
 */
#include <cuda_runtime_api.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "../../include/lc3plus_b200.h"
#include "lc3_config.h"

#define LC3B_MAGIC 0x4c433362u
#define N_PIPES 3
#define N_WORK 9
enum { K_STAGE = 0, K_FIR, K_SCALAR, K_PITCH, K_FRONT, K_SNS, K_TNS, K_RATE, K_QUANT, K_BACK, N_KERNELS };
/* kernels that read and write per-stream state: chunk c+1's launch must wait for chunk c's */
static const int k_stateful[N_KERNELS] = {1 /* input history */, 0, 1 /* high-pass memory */, 1, 0, 0, 0, 1 /* rate-loop feedback */, 0, 0};

struct LC3_Enc { /* lives in the caller's block; device state is referenced by slot */
    uint32_t magic;
    int slot[2];
    int dirty;          /* configuration words not yet mirrored to the device pool */
    int batch_mark;     /* scratch for duplicate detection inside one batch call */
    int last_bits;      /* bit depth of the frames encoded so far (0: none yet): one encoder, one bit depth */
    LC3Cfg cfg;
};

/* ------------------------------------------------------------------ device context (one per process) */
typedef struct {
    int ready, failed, device;
    cudaStream_t stream;
    /* state pool, tile-interleaved (lc3_layout.h) */
    int cap;                       /* slots, multiple of 32 */
    int32_t *cfg, *st32, *hist32;  /* device */
    int16_t *st16, *hist16;        /* device */
    int32_t *cfg_host;             /* host shadow of cfg (same layout) */
    int dirty_lo, dirty_hi;        /* tile range of cfg_host not yet uploaded */
    int *free_list, n_free, next_fresh;
    int32_t *reset_list; int n_reset, cap_reset;
    int32_t *areset_list; int n_areset, cap_areset; /* slots whose attack-detector state set_bitrate cleared (fx/setup_enc_lc3.c:266-275) */
    /* per-call lane tables: host (pinned) and device */
    size_t lanes_cap;
    int32_t *h_slot, *d_slot, *h_ostride, *d_ostride;
    int64_t *h_poff, *d_poff, *h_ooff, *d_ooff;
    /* chunk pipeline: N_PIPES chunks of frames in flight, each on its own stream with its own work buffers */
    struct {
        cudaStream_t par, ser;  /* frame-parallel kernels + copies | per-stream (stateful) kernels, higher priority */
        size_t work_cap[N_WORK]; void *work[N_WORK];
        cudaEvent_t state_ev[N_KERNELS], hop_ev[N_KERNELS + 2], done_ev;
    } pipe[N_PIPES];
    cudaEvent_t start_ev, end_ev;
    /* host-pointer calls: the chunks' input copies queue on one stream and their output copies on another, so that the
     * input of chunk c+1 streams in while chunk c computes, whatever the pipes' own streams are busy with */
    cudaStream_t h2d, d2h;
    cudaEvent_t d2h_ev;
    /* staging for host-pointer calls */
    size_t in_cap, out_cap, bfi_cap; void *d_in, *d_out, *d_bfi; uint8_t *h_bfi;
    int chunk_frames;      /* channel-frames per chunk, input resident on the device */
    int chunk_frames_host; /* same, input in host memory: shorter chunks, the copies fill and drain the pipeline faster */
    long long launches;
    /* optional per-kernel timing with CUDA events on the launching stream */
    int prof_on; cudaEvent_t *ev; int n_ev; double prof_ms[N_KERNELS + 3];
    char err[256];
} Ctx;

static Ctx g = {0};
static pthread_mutex_t g_lock = PTHREAD_MUTEX_INITIALIZER;

static int cu_fail(cudaError_t e, const char *what)
{
    if (e == cudaSuccess) return 0;
    snprintf(g.err, sizeof(g.err), "%s: %s", what, cudaGetErrorString(e));
    return 1;
}
#define CU(x) do { if (cu_fail((x), #x)) return 1; } while (0)

static size_t pool_bytes(int cap, int words, int elem) { return (size_t)cap * words * elem; }

static int pool_grow(int need)
{ /* tile-interleaved arrays grow by appending tiles, so the old contents are a prefix */
    int cap = g.cap ? g.cap : 4096;
    while (cap < need) cap *= 2;
    if (cap == g.cap) return 0;
    struct { void **p; int words, elem; } a[5] = {
        {(void **)&g.cfg, CFG_N, 4}, {(void **)&g.st32, ST32_N, 4}, {(void **)&g.hist32, LC3K_MAX_N, 4},
        {(void **)&g.st16, ST16_N, 2}, {(void **)&g.hist16, LC3K_MAX_N, 2}};
    for (int i = 0; i < 5; i++) {
        void *n = NULL;
        CU(cudaMalloc(&n, pool_bytes(cap, a[i].words, a[i].elem)));
        CU(cudaMemsetAsync(n, 0, pool_bytes(cap, a[i].words, a[i].elem), g.stream));
        if (g.cap) {
            CU(cudaMemcpyAsync(n, *a[i].p, pool_bytes(g.cap, a[i].words, a[i].elem), cudaMemcpyDeviceToDevice, g.stream));
            CU(cudaStreamSynchronize(g.stream));
            CU(cudaFree(*a[i].p));
        }
        *a[i].p = n;
    }
    int32_t *nh = (int32_t *)calloc((size_t)cap * CFG_N, 4);
    if (!nh) return 1;
    if (g.cfg_host) { memcpy(nh, g.cfg_host, pool_bytes(g.cap, CFG_N, 4)); free(g.cfg_host); }
    g.cfg_host = nh;
    int *fl = (int *)realloc(g.free_list, (size_t)cap * sizeof(int));
    if (!fl) return 1;
    g.free_list = fl;
    g.cap = cap;
    return 0;
}

static int ctx_init(void)
{
    if (g.ready) return 0;
    if (g.failed) return 1;
    g.failed = 1;
    int n = 0;
    if (cu_fail(cudaGetDeviceCount(&n), "cudaGetDeviceCount") || n <= 0) {
        if (!g.err[0]) snprintf(g.err, sizeof(g.err), "no CUDA device: liblc3plus_b200 has no CPU path");
        return 1;
    }
    CU(cudaGetDevice(&g.device));
    CU(cudaStreamCreateWithFlags(&g.stream, cudaStreamNonBlocking));
    int prio_lo = 0, prio_hi = 0;
    CU(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    for (int p = 0; p < N_PIPES; p++) {
        /* the per-stream kernels are short on parallelism (one warp or lane per stream) and sit on the critical path of
         * the next chunk: at higher priority their blocks are placed as soon as SM resources free up, so they run
         * underneath the previous chunk's frame-parallel kernels instead of queueing behind their grids */
        CU(cudaStreamCreateWithPriority(&g.pipe[p].par, cudaStreamNonBlocking, prio_lo));
        CU(cudaStreamCreateWithPriority(&g.pipe[p].ser, cudaStreamNonBlocking, prio_hi));
        for (int k = 0; k < N_KERNELS; k++) CU(cudaEventCreateWithFlags(&g.pipe[p].state_ev[k], cudaEventDisableTiming));
        for (int k = 0; k < N_KERNELS + 2; k++) CU(cudaEventCreateWithFlags(&g.pipe[p].hop_ev[k], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&g.pipe[p].done_ev, cudaEventDisableTiming));
    }
    CU(cudaEventCreate(&g.start_ev));
    CU(cudaEventCreate(&g.end_ev));
    CU(cudaStreamCreateWithFlags(&g.h2d, cudaStreamNonBlocking));
    CU(cudaStreamCreateWithFlags(&g.d2h, cudaStreamNonBlocking));
    CU(cudaEventCreateWithFlags(&g.d2h_ev, cudaEventDisableTiming));
    g.dirty_lo = 1 << 30; g.dirty_hi = -1;
    g.chunk_frames = 5 << 15;
    g.chunk_frames_host = 1 << 16;
    { const char *e = getenv("LC3PLUS_B200_HOST_CHUNK"); if (e && atoi(e) >= 32) g.chunk_frames_host = atoi(e); } /* tuning aid */
    if (pool_grow(4096)) return 1;
    g.failed = 0;
    g.ready = 1;
    return 0;
}

static int slot_alloc(void)
{
    int s;
    if (g.n_free > 0) s = g.free_list[--g.n_free];
    else {
        if (g.next_fresh >= g.cap && pool_grow(g.next_fresh + 1)) return -1;
        s = g.next_fresh++;
    }
    if (g.n_reset == g.cap_reset) {
        int nc = g.cap_reset ? g.cap_reset * 2 : 1024;
        int32_t *r = (int32_t *)realloc(g.reset_list, (size_t)nc * 4);
        if (!r) return -1;
        g.reset_list = r; g.cap_reset = nc;
    }
    g.reset_list[g.n_reset++] = s;
    return s;
}

static void cfg_to_shadow(struct LC3_Enc *e)
{
    for (int ch = 0; ch < e->cfg.channels; ch++) {
        int slot = e->slot[ch], tile = slot >> 5, lane = slot & 31;
        int32_t *w = g.cfg_host + (size_t)tile * CFG_N * 32 + lane;
        for (int k = 0; k < CFG_N; k++) w[k * 32] = e->cfg.ch[ch][k];
        if (tile < g.dirty_lo) g.dirty_lo = tile;
        if (tile > g.dirty_hi) g.dirty_hi = tile;
    }
    e->dirty = 0;
}

static int ensure(void **p, size_t *cap, size_t need, int pinned_host)
{
    if (*cap >= need) return 0;
    if (*p) { if (pinned_host) cudaFreeHost(*p); else cudaFree(*p); *p = NULL; *cap = 0; }
    size_t n = need + need / 4 + 256;
    if (pinned_host) { CU(cudaMallocHost(p, n)); } else { CU(cudaMalloc(p, n)); }
    *cap = n;
    return 0;
}

/* ------------------------------------------------------------------ batch engine */
static int is_device_ptr(const void *p)
{
    struct cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return 0; }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

static int flush_pending(void)
{ /* power-on state for freshly initialised encoders, then the configuration words */
    if (g.n_reset) {
        if (ensure((void **)&g.d_bfi, &g.bfi_cap, (size_t)g.n_reset * 4, 0)) return 1;
        CU(cudaMemcpyAsync(g.d_bfi, g.reset_list, (size_t)g.n_reset * 4, cudaMemcpyHostToDevice, g.stream));
        CU(cudaStreamSynchronize(g.stream)); /* reset_list is pageable */
        CU((cudaError_t)lc3k_launch_reset(g.st32, g.st16, g.hist16, g.hist32, (const int32_t *)g.d_bfi, g.n_reset, g.stream));
        g.launches++;
        g.n_reset = 0;
    }
    if (g.n_areset) { /* after the power-on reset: a slot may be in both lists */
        if (ensure((void **)&g.d_bfi, &g.bfi_cap, (size_t)g.n_areset * 4, 0)) return 1;
        CU(cudaMemcpyAsync(g.d_bfi, g.areset_list, (size_t)g.n_areset * 4, cudaMemcpyHostToDevice, g.stream));
        CU(cudaStreamSynchronize(g.stream)); /* areset_list is pageable */
        CU((cudaError_t)lc3k_launch_attack_reset(g.st32, (const int32_t *)g.d_bfi, g.n_areset, g.stream));
        g.launches++;
        g.n_areset = 0;
    }
    if (g.dirty_hi >= g.dirty_lo) {
        size_t off = (size_t)g.dirty_lo * CFG_N * 32, n = (size_t)(g.dirty_hi - g.dirty_lo + 1) * CFG_N * 32;
        CU(cudaMemcpyAsync(g.cfg + off, g.cfg_host + off, n * 4, cudaMemcpyHostToDevice, g.stream));
        CU(cudaStreamSynchronize(g.stream));
        g.dirty_lo = 1 << 30; g.dirty_hi = -1;
    }
    return 0;
}

static int lanes_reserve(size_t n)
{
    if (g.lanes_cap >= n) return 0;
    size_t c = n + n / 2 + 64, z;
    void **hp[4] = {(void **)&g.h_slot, (void **)&g.h_ostride, (void **)&g.h_poff, (void **)&g.h_ooff};
    void **dp[4] = {(void **)&g.d_slot, (void **)&g.d_ostride, (void **)&g.d_poff, (void **)&g.d_ooff};
    int el[4] = {4, 4, 8, 8};
    for (int i = 0; i < 4; i++) {
        z = 0; if (ensure(hp[i], &z, c * el[i], 1)) return 1;
        z = 0; if (ensure(dp[i], &z, c * el[i], 0)) return 1;
    }
    g.lanes_cap = c;
    return 0;
}

/* Frames of chunk c when `rem` frames are left and a full chunk is fc frames.  With host input (ramp) the first two chunks
 * are a quarter and a half chunk, so that the kernels start after a short first copy instead of a full chunk's. */
static int chunk_size(int c, int rem, int fc, int ramp)
{
    int sz;
    if (ramp && c == 0) sz = fc / 4;
    else if (ramp && c == 1) sz = fc / 2;
    else { int k = (rem + fc - 1) / fc; sz = (rem + k - 1) / k; }
    if (sz < 1) sz = 1;
    return sz < rem ? sz : rem;
}

/* one bucket: encoders that share (fs_idx, dms).  idx[] lists their positions in encs[]. */
typedef struct { const void *h_pcm; void *h_out; } ChunkCopy; /* host buffers to copy chunk by chunk (NULL: not needed) */

static int run_bucket(struct LC3_Enc *const *encs, const int *idx, int n, int n_frames, const void *d_pcm, size_t stride,
                      int bits, uint8_t *d_out, size_t ostride, const int *num_bytes_in, const ChunkCopy *cp)
{
    const LC3Cfg *c0 = &encs[idx[0]]->cfg;
    const int N = c0->N, sb = bits == 16 ? 2 : 4;
    size_t n_lanes = 0;
    for (int i = 0; i < n; i++) n_lanes += encs[idx[i]]->cfg.channels;
    if (lanes_reserve(n_lanes)) return 1;
    int any_bfi = 0, max_bytes = 0, any_ep = 0;
    size_t l = 0;
    for (int i = 0; i < n; i++) {
        const struct LC3_Enc *e = encs[idx[i]];
        int nb = lc3cfg_num_bytes(&e->cfg), boff = 0;
        any_ep |= e->cfg.epmode;
        for (int ch = 0; ch < e->cfg.channels; ch++, l++) {
            if (e->cfg.ch[ch][CFG_TARGET_BYTES] > max_bytes) max_bytes = e->cfg.ch[ch][CFG_TARGET_BYTES];
            g.h_slot[l] = e->slot[ch];
            g.h_poff[l] = (int64_t)((size_t)idx[i] * stride / sb) + (int64_t)ch * n_frames * N;
            g.h_ooff[l] = (int64_t)((size_t)idx[i] * ostride) + boff;
            g.h_ostride[l] = nb;
            boff += lc3cfg_out_bytes(&e->cfg, ch);
        }
        if (num_bytes_in)
            for (int f = 0; f < n_frames && !any_bfi; f++) any_bfi = num_bytes_in[(size_t)idx[i] * n_frames + f] == -1;
    }
    CU(cudaMemcpyAsync(g.d_slot, g.h_slot, n_lanes * 4, cudaMemcpyHostToDevice, g.stream));
    CU(cudaMemcpyAsync(g.d_ostride, g.h_ostride, n_lanes * 4, cudaMemcpyHostToDevice, g.stream));
    CU(cudaMemcpyAsync(g.d_poff, g.h_poff, n_lanes * 8, cudaMemcpyHostToDevice, g.stream));
    CU(cudaMemcpyAsync(g.d_ooff, g.h_ooff, n_lanes * 8, cudaMemcpyHostToDevice, g.stream));

    LC3K_Batch b;
    memset(&b, 0, sizeof(b));
    lc3cfg_fill_batch(c0, &b);
    b.bits = bits;
    b.n_lanes = (int)n_lanes;
    b.n_tiles = (int)((n_lanes + 31) / 32);
    b.max_bytes = max_bytes;
    b.slot = g.d_slot; b.cfg = g.cfg; b.st32 = g.st32; b.st16 = g.st16; b.hist16 = g.hist16; b.hist32 = g.hist32;
    b.pcm = d_pcm; b.pcm_off = g.d_poff; b.out = d_out; b.out_off = g.d_ooff; b.out_stride = g.d_ostride;

    if (any_bfi) { /* [n_frames][n_lanes] flags */
        size_t nbf = (size_t)n_frames * n_lanes, z = 0;
        uint8_t *hb = NULL;
        if (ensure((void **)&hb, &z, nbf, 1)) return 1;
        l = 0;
        for (int i = 0; i < n; i++)
            for (int ch = 0; ch < encs[idx[i]]->cfg.channels; ch++, l++)
                for (int f = 0; f < n_frames; f++) hb[(size_t)f * n_lanes + l] = num_bytes_in[(size_t)idx[i] * n_frames + f] == -1;
        if (ensure(&g.d_bfi, &g.bfi_cap, nbf, 0)) { cudaFreeHost(hb); return 1; }
        cudaError_t e1 = cudaMemcpyAsync(g.d_bfi, hb, nbf, cudaMemcpyHostToDevice, g.stream);
        cudaError_t e2 = cudaStreamSynchronize(g.stream);
        cudaFreeHost(hb);
        CU(e1); CU(e2);
        b.bfi = (const uint8_t *)g.d_bfi;
    }

    /* frames per chunk: about g.chunk_frames channel-frames each, balanced over the chunks */
    int fc = ((cp && cp->h_pcm) ? g.chunk_frames_host : g.chunk_frames) / (b.n_tiles * 32);
    if (fc < 1) fc = 1;
    if (fc > n_frames) fc = n_frames;
    const int ramp = cp && cp->h_pcm && n_frames > fc;
    int n_chunks = 0;
    for (int rem = n_frames; rem > 0; n_chunks++) rem -= chunk_size(n_chunks, rem, fc, ramp);
    const int n_pipes = n_chunks < N_PIPES ? n_chunks : N_PIPES;
    const size_t ftl = (size_t)fc * b.n_tiles * 32;
    const size_t need[N_WORK] = {(size_t)b.n_tiles * 32 * (size_t)(fc + 1) * N * sb, ftl * PO_N * 4, ftl * N * 4, ftl * FO_N * 4,
                                 ftl * (N / 4) * 4, ftl * RO_N * 4, ftl * 128 * 2, ftl * N * 2, ftl * BEN_N * 4};
    for (int p = 0; p < n_pipes; p++)
        for (int i = 0; i < N_WORK; i++)
            if (ensure(&g.pipe[p].work[i], &g.pipe[p].work_cap[i], need[i], 0)) return 1;

    const int EPC = 2 * N_KERNELS; /* profiling events per chunk: one before and one after every kernel */
    if (g.prof_on && g.n_ev < n_chunks * EPC) {
        cudaEvent_t *ne = (cudaEvent_t *)realloc(g.ev, (size_t)n_chunks * EPC * sizeof(cudaEvent_t));
        if (!ne) return 1;
        g.ev = ne;
        for (; g.n_ev < n_chunks * EPC; g.n_ev++) CU(cudaEventCreate(&g.ev[g.n_ev]));
    }
    int (*const launch[N_KERNELS])(const LC3K_Batch *, void *) = {lc3k_launch_stage, lc3k_launch_fir, lc3k_launch_scalar, lc3k_launch_pitch,
                                                                  lc3k_launch_front, lc3k_launch_sns, lc3k_launch_tns, lc3k_launch_rate, lc3k_launch_quant, lc3k_launch_back};
    /* everything queued on the main stream so far (state reset, configuration, lane tables, whole-buffer input copy)
     * is ordered before the first chunk of every pipe */
    CU(cudaEventRecord(g.start_ev, g.stream));
    for (int f0 = 0, c = 0, cf; f0 < n_frames; f0 += cf, c++) {
        cf = chunk_size(c, n_frames - f0, fc, ramp);
        const int p = c % N_PIPES;
        const cudaStream_t par = g.pipe[p].par, ser = g.pipe[p].ser;
        void *const *w = g.pipe[p].work;
        if (c < N_PIPES) { CU(cudaStreamWaitEvent(par, g.start_ev, 0)); CU(cudaStreamWaitEvent(ser, g.start_ev, 0)); }
        else CU(cudaStreamWaitEvent(ser, g.pipe[p].done_ev, 0)); /* chunk c - N_PIPES is done with these work buffers */
        b.frame0 = f0;
        b.n_frames = cf;
        b.pcm_t = w[0]; b.pout = (int32_t *)w[1]; b.spec = (int32_t *)w[2]; b.fout = (int32_t *)w[3];
        b.en4 = (int32_t *)w[4]; b.rout = (int32_t *)w[5]; b.r12 = (int16_t *)w[6]; b.xq = (int16_t *)w[7]; b.ben = (int32_t *)w[8];
        const int copy_in = cp && cp->h_pcm;
        if (copy_in) { /* this chunk's columns of every stream's PCM rows, one rectangle per channel */
            if (c == 0) CU(cudaStreamWaitEvent(g.h2d, g.start_ev, 0));
            for (int ch = 0; ch < c0->channels; ch++) {
                const size_t off = ((size_t)ch * n_frames + f0) * N * sb;
                CU(cudaMemcpy2DAsync((char *)d_pcm + off, stride, (const char *)cp->h_pcm + off, stride, (size_t)b.n_frames * N * sb,
                                     (size_t)n, cudaMemcpyHostToDevice, g.h2d));
            }
            CU(cudaEventRecord(g.pipe[p].hop_ev[N_KERNELS], g.h2d));
        }
        /* The chunk's kernels form a small graph over the two streams.  16-bit input, long frames: the tile MDCT needs only
         * the staged samples, so it is issued right behind the FIR on `par` and runs beside the per-stream high-pass and
         * pitch passes on `ser`; SNS then waits for both (it takes the attack flag from the pitch pass).  Otherwise (24-bit
         * input: the block exponent comes from the pitch pass; short frames: one fused front-end kernel) it is a chain. */
        const int split = bits == 16 && N >= 160 && !getenv("LC3PLUS_B200_CHAIN"); /* the variable forces the chain: per-kernel timing (tools/kernel_ms.py) */
        static const int order_chain[N_KERNELS] = {K_STAGE, K_FIR, K_SCALAR, K_PITCH, K_FRONT, K_SNS, K_TNS, K_RATE, K_QUANT, K_BACK};
        static const int order_split[N_KERNELS] = {K_STAGE, K_FIR, K_FRONT, K_SCALAR, K_PITCH, K_SNS, K_TNS, K_RATE, K_QUANT, K_BACK};
        /* producers on the OTHER stream that kernel k must wait for (same-stream producers are ordered by the stream) */
        static const int dep_chain[N_KERNELS][2] = {{-1, -1}, {K_STAGE, -1}, {K_FIR, -1}, {-1, -1}, {K_PITCH, -1}, {-1, -1}, {-1, -1},
                                                    {K_TNS, K_FRONT}, {K_RATE, -1}, {-1, -1}};
        static const int dep_split[N_KERNELS][2] = {{-1, -1}, {K_STAGE, -1}, {K_FIR, -1}, {-1, -1}, {-1, -1}, {K_PITCH, -1}, {-1, -1},
                                                    {K_TNS, -1}, {K_RATE, -1}, {-1, -1}};
        const int *order = split ? order_split : order_chain;
        int ran[N_KERNELS] = {0};
        cudaStream_t first = NULL;
        for (int o = 0; o < N_KERNELS; o++) {
            const int k = order[o];
            if (bits != 16 && (k == K_FIR || k == K_SCALAR)) continue; /* fused into K_PITCH for 24-bit input */
            if (N < 160 && (k == K_SNS || k == K_TNS)) continue;        /* fused into K_FRONT for the short frames */
            const cudaStream_t st = k_stateful[k] ? ser : par;
            if (!first) { /* the first kernel (K_STAGE) waits for the chunk's input copy */
                first = st;
                if (copy_in) CU(cudaStreamWaitEvent(st, g.pipe[p].hop_ev[N_KERNELS], 0));
            }
            const int *dep = split ? dep_split[k] : dep_chain[k];
            for (int j = 0; j < 2; j++)
                if (dep[j] >= 0 && ran[dep[j]] && (k_stateful[dep[j]] ? ser : par) != st) CU(cudaStreamWaitEvent(st, g.pipe[p].hop_ev[dep[j]], 0));
            if (k_stateful[k] && c > 0) CU(cudaStreamWaitEvent(st, g.pipe[(c - 1) % N_PIPES].state_ev[k], 0));
            if (g.prof_on) CU(cudaEventRecord(g.ev[c * EPC + 2 * k], st));
            CU((cudaError_t)launch[k](&b, st));
            if (g.prof_on) CU(cudaEventRecord(g.ev[c * EPC + 2 * k + 1], st));
            CU(cudaEventRecord(g.pipe[p].hop_ev[k], st));
            ran[k] = 1;
            if (k_stateful[k] && c + 1 < n_chunks) CU(cudaEventRecord(g.pipe[p].state_ev[k], st));
        }
        /* K_BACK ran on `par`; error protection, if any lane asks for it, codes the finished payloads in place */
        if (any_ep) { CU((cudaError_t)lc3k_launch_fec(&b, par)); g.launches++; }
        CU(cudaEventRecord(g.pipe[p].done_ev, par));
        if (cp && cp->h_out) {
            const size_t nb = (size_t)lc3cfg_num_bytes(c0), off = (size_t)f0 * nb;
            CU(cudaStreamWaitEvent(g.d2h, g.pipe[p].done_ev, 0));
            CU(cudaMemcpy2DAsync((char *)cp->h_out + off, ostride, d_out + off, ostride, (size_t)b.n_frames * nb, (size_t)n,
                                 cudaMemcpyDeviceToHost, g.d2h));
        }
        g.launches += (bits == 16 ? 8 : 6) + (N >= 160 ? 2 : 0); /* frame lengths >= 160: the front end is three kernels (tile MDCT, SNS, TNS) */
    }
    for (int p = 0; p < n_pipes; p++) CU(cudaStreamWaitEvent(g.stream, g.pipe[p].done_ev, 0));
    if (cp && cp->h_out) { CU(cudaEventRecord(g.d2h_ev, g.d2h)); CU(cudaStreamWaitEvent(g.stream, g.d2h_ev, 0)); }
    CU(cudaEventRecord(g.end_ev, g.stream));
    CU(cudaStreamSynchronize(g.stream)); /* lane tables are reused by the next bucket */
    if (g.prof_on) {
        /* per-kernel time = elapsed on the kernel's own stream; with several chunks in flight the kernels of different
         * chunks share the SMs, so these sum to more than the span */
        float ms;
        for (int c = 0; c < n_chunks; c++)
            for (int k = 0; k < N_KERNELS; k++) {
                if ((bits != 16 && (k == K_FIR || k == K_SCALAR)) || (N < 160 && (k == K_SNS || k == K_TNS))) continue;
                CU(cudaEventElapsedTime(&ms, g.ev[c * EPC + 2 * k], g.ev[c * EPC + 2 * k + 1]));
                g.prof_ms[k] += ms;
            }
        CU(cudaEventElapsedTime(&ms, g.start_ev, g.end_ev));
        g.prof_ms[N_KERNELS] += ms;
        g.prof_ms[N_KERNELS + 1] += 1;
        g.prof_ms[N_KERNELS + 2] += n_chunks;
    }
    return 0;
}

static unsigned g_call_id = 0;

static LC3_Error run_batch(struct LC3_Enc *const *encs, int n_streams, int n_frames, const void *pcm, size_t stride, int bits,
                           void *out, size_t ostride, int *num_bytes)
{
    if (!encs || !pcm || !out) return LC3_NULL_ERROR;
    if (n_streams <= 0 || n_frames <= 0 || (bits != 16 && bits != 24)) return LC3_ERROR;
    const int sb = bits == 16 ? 2 : 4;
    for (int i = 0; i < n_streams; i++) {
        if (!encs[i]) return LC3_NULL_ERROR;
        if (encs[i]->magic != LC3B_MAGIC) return LC3_ERROR;
        if (!encs[i]->cfg.br_set) return LC3_BITRATE_UNSET_ERROR;
        if (encs[i]->last_bits && encs[i]->last_bits != bits) {
            /* the reference carries the MDCT overlap across an lc3_enc16 <-> lc3_enc24 switch (fx/enc_lc3.c:35-83); here the two
             * input widths keep their history in different pools, so the switch is refused rather than coded differently */
            snprintf(g.err, sizeof(g.err), "encoder %d: bit depth changed from %d to %d (one encoder, one bit depth: re-initialise it)", i, encs[i]->last_bits, bits);
            return LC3_ERROR;
        }
    }
    pthread_mutex_lock(&g_lock);
    LC3_Error ret = LC3_ERROR;
    int *order = NULL;
    if (ctx_init()) goto done;
    cudaSetDevice(g.device);
    unsigned id = ++g_call_id;
    size_t max_in = 0, max_out = 0;
    int uniform = 1, one_bucket = 1;
    for (int i = 0; i < n_streams; i++) {
        struct LC3_Enc *e = encs[i];
        if (e->cfg.dms_idx != encs[0]->cfg.dms_idx || e->cfg.fs_idx != encs[0]->cfg.fs_idx || e->cfg.channels != encs[0]->cfg.channels)
            one_bucket = 0;
        if ((unsigned)e->batch_mark == id) { snprintf(g.err, sizeof(g.err), "encoder listed twice in one batch"); goto done; }
        e->batch_mark = (int)id;
        if (e->dirty) cfg_to_shadow(e);
        size_t in_b = (size_t)e->cfg.channels * n_frames * e->cfg.N * sb, out_b = (size_t)n_frames * lc3cfg_num_bytes(&e->cfg);
        if (i && (in_b != max_in || out_b != max_out)) uniform = 0;
        if (in_b > max_in) max_in = in_b;
        if (out_b > max_out) max_out = out_b;
        if (n_streams > 1 && (stride < in_b || ostride < out_b)) { snprintf(g.err, sizeof(g.err), "stream stride too small"); goto done; }
    }
    if (flush_pending()) goto done;

    const int pcm_dev = is_device_ptr(pcm), out_dev = is_device_ptr(out);
    const void *d_pcm = pcm;
    uint8_t *d_out = (uint8_t *)out;
    /* equal-sized streams of one (sample rate, frame size): host buffers are copied chunk by chunk on the chunk's own
     * stream, so the copies of one chunk overlap the kernels of its neighbours; otherwise whole-buffer copies up front */
    const int chunked = uniform && one_bucket;
    ChunkCopy cp = {chunked && !pcm_dev ? pcm : NULL, chunked && !out_dev ? out : NULL};
    if (!pcm_dev) {
        if (ensure(&g.d_in, &g.in_cap, (size_t)n_streams * stride + max_in, 0)) goto done;
        if (chunked) { /* copied per chunk in run_bucket */
        } else if (uniform) {
            if (cu_fail(cudaMemcpy2DAsync(g.d_in, stride, pcm, stride, max_in, n_streams, cudaMemcpyHostToDevice, g.stream), "H2D pcm")) goto done;
        } else {
            for (int i = 0; i < n_streams; i++) {
                size_t in_b = (size_t)encs[i]->cfg.channels * n_frames * encs[i]->cfg.N * sb;
                if (cu_fail(cudaMemcpyAsync((char *)g.d_in + i * stride, (const char *)pcm + i * stride, in_b, cudaMemcpyHostToDevice, g.stream), "H2D pcm")) goto done;
            }
        }
        d_pcm = g.d_in;
    }
    if (!out_dev) {
        if (ensure(&g.d_out, &g.out_cap, (size_t)n_streams * ostride + max_out, 0)) goto done;
        d_out = (uint8_t *)g.d_out;
    }

    /* bucket by (fs_idx, dms_idx) */
    order = (int *)malloc((size_t)n_streams * sizeof(int));
    if (!order) goto done;
    for (int key = 0; key < 15; key++) {
        int n = 0;
        for (int i = 0; i < n_streams; i++)
            if (encs[i]->cfg.dms_idx * 5 + encs[i]->cfg.fs_idx == key) order[n++] = i;
        if (n && run_bucket(encs, order, n, n_frames, d_pcm, stride, bits, d_out, ostride, num_bytes, &cp)) goto done;
    }
    if (!out_dev && !chunked) {
        if (uniform) {
            if (cu_fail(cudaMemcpy2DAsync(out, ostride, d_out, ostride, max_out, n_streams, cudaMemcpyDeviceToHost, g.stream), "D2H out")) goto done;
        } else {
            for (int i = 0; i < n_streams; i++) {
                size_t out_b = (size_t)n_frames * lc3cfg_num_bytes(&encs[i]->cfg);
                if (cu_fail(cudaMemcpyAsync((char *)out + i * ostride, d_out + i * ostride, out_b, cudaMemcpyDeviceToHost, g.stream), "D2H out")) goto done;
            }
        }
    }
    if (cu_fail(cudaStreamSynchronize(g.stream), "sync")) goto done;
    for (int i = 0; i < n_streams; i++) encs[i]->last_bits = bits;
    if (num_bytes)
        for (int i = 0; i < n_streams; i++) {
            int nb = lc3cfg_num_bytes(&encs[i]->cfg);
            for (int f = 0; f < n_frames; f++) num_bytes[(size_t)i * n_frames + f] = nb;
        }
    ret = LC3_OK;
done:
    free(order);
    pthread_mutex_unlock(&g_lock);
    return ret;
}

/* ------------------------------------------------------------------ reference API (fx/lc3.h, fx/lc3.c) */
int lc3_version(void) { return LC3_VERSION; }                                   /* fx/lc3.c:30-33 */
int lc3_channels_supported(int channels) { return channels >= 1 && channels <= LC3_MAX_CHANNELS; } /* fx/lc3.c:35-38 */
int lc3_samplerate_supported(int samplerate)
{ /* fx/lc3.c:40-54 */
    switch (samplerate) {
    case 8000: case 16000: case 24000: case 32000: case 44100: case 48000: return 1;
    default: return 0;
    }
}

int lc3_enc_get_size(int samplerate, int channels)
{ /* fx/lc3.c:133-138: 0 for unsupported arguments */
    if (!lc3_samplerate_supported(samplerate) || !lc3_channels_supported(channels)) return 0;
    return (int)sizeof(struct LC3_Enc);
}

LC3_Error lc3_enc_init(LC3_Enc *encoder, int samplerate, int channels)
{ /* fx/lc3.c:124-131: NULL -> alignment -> samplerate -> channels */
    if (encoder == NULL) return LC3_NULL_ERROR;
    if ((uintptr_t)encoder % 4 != 0) return LC3_ALIGN_ERROR;
    if (!lc3_samplerate_supported(samplerate)) return LC3_SAMPLERATE_ERROR;
    if (!lc3_channels_supported(channels)) return LC3_CHANNELS_ERROR;
    memset(encoder, 0, sizeof(*encoder));
    int r = lc3cfg_init(&encoder->cfg, samplerate, channels);
    if (r) return (LC3_Error)r;
    pthread_mutex_lock(&g_lock);
    int ok = !ctx_init();
    if (ok) cudaSetDevice(g.device);
    for (int ch = 0; ok && ch < channels; ch++) {
        encoder->slot[ch] = slot_alloc();
        if (encoder->slot[ch] < 0) ok = 0;
    }
    pthread_mutex_unlock(&g_lock);
    if (!ok) return LC3_ERROR; /* no device: there is no CPU path */
    encoder->magic = LC3B_MAGIC;
    return LC3_OK;
}

void lc3plus_enc_release(LC3_Enc *encoder)
{
    if (!encoder || encoder->magic != LC3B_MAGIC) return;
    pthread_mutex_lock(&g_lock);
    for (int ch = 0; ch < encoder->cfg.channels; ch++) g.free_list[g.n_free++] = encoder->slot[ch];
    pthread_mutex_unlock(&g_lock);
    encoder->magic = 0;
}

static int enc_ok(const LC3_Enc *e) { return e != NULL && e->magic == LC3B_MAGIC; }

int lc3_enc_get_scratch_size(const LC3_Enc *encoder)
{ /* fx/lc3.c:140-150 (the scratch is not used by this implementation; the size is kept for callers) */
    if (!enc_ok(encoder)) return 0;
    int n = encoder->cfg.N > 160 ? encoder->cfg.N : 160;
    return 14 * n + 64;
}
int lc3_enc_get_input_samples(const LC3_Enc *encoder) { return enc_ok(encoder) ? encoder->cfg.N : 0; }       /* fx/lc3.c:152-156 */
int lc3_enc_get_num_bytes(const LC3_Enc *encoder) { return enc_ok(encoder) ? lc3cfg_num_bytes(&encoder->cfg) : 0; } /* fx/lc3.c:158-161 */
int lc3_enc_get_real_bitrate(const LC3_Enc *encoder) { return enc_ok(encoder) ? lc3cfg_real_bitrate(&encoder->cfg) : 0; } /* fx/lc3.c:163-184 */
int lc3_enc_get_delay(const LC3_Enc *encoder)
{ /* fx/lc3.c:193-197 */
    return enc_ok(encoder) ? encoder->cfg.N - 2 * encoder->cfg.la_zeroes : 0;
}

LC3_Error lc3_enc_set_bitrate(LC3_Enc *encoder, int bitrate)
{ /* fx/lc3.c:186-191 */
    if (encoder == NULL) return LC3_NULL_ERROR;
    if (encoder->magic != LC3B_MAGIC) return LC3_ERROR;
    int r = lc3cfg_set_bitrate(&encoder->cfg, bitrate);
    if (r == 0) {
        encoder->dirty = 1;
        /* the reference clears the attack detector's state right here whenever the new bitrate switches it off, frames in
         * between or not (fx/setup_enc_lc3.c:266-275): queue the slot for a clear before the next batch */
        pthread_mutex_lock(&g_lock);
        for (int ch = 0; ch < encoder->cfg.channels; ch++) {
            if (!encoder->cfg.attack_reset[ch]) continue;
            encoder->cfg.attack_reset[ch] = 0;
            if (g.n_areset == g.cap_areset) {
                int nc = g.cap_areset ? g.cap_areset * 2 : 1024;
                int32_t *a = (int32_t *)realloc(g.areset_list, (size_t)nc * 4);
                if (!a) { pthread_mutex_unlock(&g_lock); return LC3_ERROR; }
                g.areset_list = a; g.cap_areset = nc;
            }
            g.areset_list[g.n_areset++] = encoder->slot[ch];
        }
        pthread_mutex_unlock(&g_lock);
    }
    return (LC3_Error)r;
}
LC3_Error lc3_enc_set_frame_dms(LC3_Enc *encoder, int frame_ms)
{ /* fx/lc3.c:215-223 */
    if (encoder == NULL) return LC3_NULL_ERROR;
    if (encoder->magic != LC3B_MAGIC) return LC3_ERROR;
    return (LC3_Error)lc3cfg_set_frame_dms(&encoder->cfg, frame_ms);
}
LC3_Error lc3_enc_set_bandwidth(LC3_Enc *encoder, int bandwidth)
{ /* fx/lc3.c:226-244 */
    if (encoder == NULL) return LC3_NULL_ERROR;
    if (encoder->magic != LC3B_MAGIC) return LC3_ERROR;
    int r = lc3cfg_set_bandwidth(&encoder->cfg, bandwidth);
    if (r == 0) encoder->dirty = 1;
    return (LC3_Error)r;
}
LC3_Error lc3_enc_set_ep_mode(LC3_Enc *encoder, LC3_EpMode epmode)
{ /* fx/lc3.c:199-205 */
    if (encoder == NULL) return LC3_NULL_ERROR;
    if ((unsigned)epmode > LC3_EP_HIGH) return LC3_EPMODE_ERROR;
    if (encoder->magic != LC3B_MAGIC) return LC3_ERROR;
    int r = lc3cfg_set_ep_mode(&encoder->cfg, (int)epmode);
    if (r == 0) encoder->dirty = 1;
    return (LC3_Error)r;
}
LC3_Error lc3_enc_set_ep_mode_request(LC3_Enc *encoder, LC3_EpModeRequest epmr)
{ /* fx/lc3.c:207-213; the request is carried in error-protected frames */
    if (encoder == NULL) return LC3_NULL_ERROR;
    if ((unsigned)epmr > LC3_EPMR_HIGH) return LC3_EPMR_ERROR;
    if (encoder->magic != LC3B_MAGIC) return LC3_ERROR;
    int r = lc3cfg_set_ep_mode_request(&encoder->cfg, (int)epmr);
    if (r == 0) encoder->dirty = 1;
    return (LC3_Error)r;
}

/* one frame through the batch engine: gather the caller's per-channel pointers into a planar block */
static LC3_Error enc_one(LC3_Enc *encoder, void **in, int bits, void *out, int *num_bytes)
{ /* fx/lc3.c:246-257 */
    if (encoder == NULL || in == NULL || out == NULL || num_bytes == NULL) return LC3_NULL_ERROR;
    if (encoder->magic != LC3B_MAGIC) return LC3_ERROR;
    for (int ch = 0; ch < encoder->cfg.channels; ch++)
        if (in[ch] == NULL) return LC3_NULL_ERROR;
    if (!encoder->cfg.br_set) return LC3_BITRATE_UNSET_ERROR;
    int32_t planar[2 * LC3K_MAX_N];
    const int N = encoder->cfg.N, sb = bits == 16 ? 2 : 4;
    for (int ch = 0; ch < encoder->cfg.channels; ch++) memcpy((char *)planar + (size_t)ch * N * sb, in[ch], (size_t)N * sb);
    uint8_t frame[LC3_MAX_BYTES];
    LC3_Enc *list[1] = {encoder};
    int nb = *num_bytes;
    LC3_Error r = run_batch(list, 1, 1, planar, sizeof(planar), bits, frame, sizeof(frame), &nb);
    if (r != LC3_OK) return r;
    memcpy(out, frame, (size_t)nb);
    *num_bytes = nb;
    return LC3_OK;
}
LC3_Error lc3_enc16(LC3_Enc *encoder, int16_t **input_samples, void *output_bytes, int *num_bytes, void *scratch)
{ /* fx/lc3.c:259-262 */
    (void)scratch;
    return enc_one(encoder, (void **)input_samples, 16, output_bytes, num_bytes);
}
LC3_Error lc3_enc24(LC3_Enc *encoder, int32_t **input_samples, void *output_bytes, int *num_bytes, void *scratch)
{ /* fx/lc3.c:264-267 */
    (void)scratch;
    return enc_one(encoder, (void **)input_samples, 24, output_bytes, num_bytes);
}

LC3_Error lc3plus_enc_batch(LC3_Enc *const *encoders, int n_streams, int n_frames, const void *pcm, size_t stream_stride_bytes,
                            int bitdepth, void *out, size_t out_stream_stride_bytes, int *num_bytes)
{
    return run_batch(encoders, n_streams, n_frames, pcm, stream_stride_bytes, bitdepth, out, out_stream_stride_bytes, num_bytes);
}

long long lc3plus_kernel_launches(void) { return g.launches; }
void lc3plus_profile_enable(int on) { g.prof_on = on; memset(g.prof_ms, 0, sizeof(g.prof_ms)); }
void lc3plus_profile_read(double *ms) { memcpy(ms, g.prof_ms, sizeof(g.prof_ms)); }
const char *lc3plus_last_error(void) { return g.err; }
void lc3b_set_error(const char *msg) { snprintf(g.err, sizeof(g.err), "%s", msg); } /* for lc3_k_noalloc.cu */
void lc3plus_set_chunk_frames(int channel_frames)
{ /* >= 32: both chunk sizes; 0: back to the defaults */
    pthread_mutex_lock(&g_lock);
    if (!ctx_init()) {
        if (channel_frames >= 32) g.chunk_frames = g.chunk_frames_host = channel_frames;
        else if (channel_frames == 0) { g.chunk_frames = 5 << 15; g.chunk_frames_host = 1 << 16; }
    }
    pthread_mutex_unlock(&g_lock);
}

/* `lc3plus_` spellings */
LC3_Error lc3plus_enc_init(LC3_Enc *e, int fs, int ch) { return lc3_enc_init(e, fs, ch); }
LC3_Error lc3plus_enc16(LC3_Enc *e, int16_t **in, void *out, int *nb, void *s) { return lc3_enc16(e, in, out, nb, s); }
LC3_Error lc3plus_enc24(LC3_Enc *e, int32_t **in, void *out, int *nb, void *s) { return lc3_enc24(e, in, out, nb, s); }
LC3_Error lc3plus_enc_set_bitrate(LC3_Enc *e, int br) { return lc3_enc_set_bitrate(e, br); }
LC3_Error lc3plus_enc_set_frame_dms(LC3_Enc *e, int dms) { return lc3_enc_set_frame_dms(e, dms); }
LC3_Error lc3plus_enc_set_bandwidth(LC3_Enc *e, int bw) { return lc3_enc_set_bandwidth(e, bw); }
int lc3plus_enc_get_size(int fs, int ch) { return lc3_enc_get_size(fs, ch); }
int lc3plus_enc_get_input_samples(const LC3_Enc *e) { return lc3_enc_get_input_samples(e); }
int lc3plus_enc_get_num_bytes(const LC3_Enc *e) { return lc3_enc_get_num_bytes(e); }
int lc3plus_enc_get_real_bitrate(const LC3_Enc *e) { return lc3_enc_get_real_bitrate(e); }
int lc3plus_enc_get_delay(const LC3_Enc *e) { return lc3_enc_get_delay(e); }
