/* Host C layer of liblc3plus_b200: the reference encoder API (fx/lc3.h, fx/lc3.c:94-267) and the batched entry points on
 * top of the CUDA kernels.  Plain C; talks to the device through the CUDA runtime C API and the lc3k_launch_* functions.
 * There is no CPU encode path in this library.
 *
 * One context PER DEVICE (state pool, streams, work buffers, its own lock).  An encoder lives on the device that was
 * selected (lc3plus_set_device; default: the calling thread's current CUDA device) when lc3_enc_init ran; a batch whose
 * encoders live on several devices is queued by one host thread per device and joined at the end.
 * The chunks of frames form ONE continuing pipeline per device: the chunks of a call follow the previous call's without
 * draining the device, so with lc3plus_enc_batch_async the copies and kernels of consecutive calls overlap. */
#include <cuda_runtime_api.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "../../include/lc3plus_b200.h"
#include "lc3_config.h"

#define LC3B_MAGIC 0x4c433362u
#define LC3B_MAX_DEV 16
#define N_PIPES 8      /* pipes that exist */
#define N_PIPES_MIN 3  /* pipes in use: three chunks in flight (four measured slower on one bucket), more when a call has
                          many small (sample rate, frame size) buckets, whose warp-per-stream kernels fill a fraction of the device */
#define N_WORK 9
#define N_BUCKETS 15
enum { K_STAGE = 0, K_FIR, K_SCALAR, K_PITCH, K_FRONT, K_SNS, K_TNS, K_RATE, K_QUANT, K_BACK, N_KERNELS };
/* kernels that read and write per-stream state: chunk c+1's launch must wait for chunk c's */
static const int k_stateful[N_KERNELS] = {1 /* input history */, 0, 1 /* high-pass memory */, 1, 0, 0, 0, 1 /* rate-loop feedback */, 0, 0};

struct LC3_Enc { /* lives in the caller's block; device state is referenced by slot */
    uint32_t magic;
    int device;         /* CUDA device that holds the state */
    int slot[2];
    int dirty;          /* configuration words not yet mirrored to the device pool */
    int batch_mark;     /* scratch for duplicate detection inside one batch call */
    int last_bits;      /* bit depth of the frames encoded so far (0: none yet): one encoder, one bit depth */
    LC3Cfg cfg;
};

/* lane tables of one bucket, one block: slot[n] i32 | ostride[n] i32 | poff[n] i64 | ooff[n] i64.  Kept from call to call:
 * a call that finds the same table (the usual streaming case) uploads nothing. */
typedef struct { size_t cap, n; unsigned char *h /* pinned */, *d[2]; int cur; } Tab; /* two device copies: a changed table goes
   into the one no queued kernel reads (the call before the previous one has finished), like the configuration words */

typedef struct {
    int device;
    pthread_mutex_t lock;
    cudaStream_t up;               /* uploads: resets, configuration, lane tables, whole-buffer input copies */
    cudaStream_t join;             /* joins the end of every call; whole-buffer output copies */
    /* state pool, tile-interleaved (lc3_layout.h) */
    int cap;                       /* slots, multiple of 32 */
    int32_t *cfg, *cfg_b, *st32, *hist32;  /* device; cfg / cfg_b: the configuration words alternate between two copies, so new
                                      values can be uploaded while queued kernels still read the old ones */
    int cfg_cur;                   /* 0: kernels of the current call read cfg, 1: cfg_b */
    int16_t *st16, *hist16;        /* device */
    int32_t *cfg_host;             /* pinned host shadow of the configuration words (same layout) */
    cudaEvent_t cfg_ev; int cfg_ev_used; /* the last upload from the shadow has been read */
    cudaEvent_t tab_ev; int tab_ev_used; /* the last lane-table upload has read its pinned source */
    int dirty_lo, dirty_hi;        /* tile range of cfg_host not yet uploaded */
    int *free_list, n_free, next_fresh;
    int32_t *reset_list; int n_reset, cap_reset;    /* slots to set to their power-on state */
    void *d_list; size_t list_cap;
    Tab tab[N_BUCKETS];
    unsigned char *tab_tmp; size_t tab_tmp_cap;
    /* chunk pipeline: N_PIPES chunks of frames in flight, each on its own streams with its own work buffers */
    struct {
        cudaStream_t par, ser;  /* frame-parallel kernels | per-stream (stateful) kernels, higher priority */
        size_t work_cap[N_WORK]; void *work[N_WORK];
        cudaEvent_t state_ev[N_KERNELS], hop_ev[N_KERNELS + 2], done_ev;
        int used;               /* done_ev / state_ev have been recorded at least once */
    } pipe[N_PIPES];
    long long seq;              /* chunks queued so far (runs on across calls) */
    int n_active, next_pipe, last_pipe; /* pipes in rotation for this call, the next one to use, the previous chunk's (-1: none) */
    cudaEvent_t call_ev;        /* this call's uploads are done */
    cudaEvent_t call_end_ev[2]; int call_end_used[2]; long long call_no;
    /* host-pointer calls: input copies queue on one stream, output copies on another */
    cudaStream_t h2d, d2h;
    cudaEvent_t d2h_ev;
    size_t in_cap[2], out_cap[2]; void *d_in[2], *d_out[2];   /* staging, alternating between calls */
    size_t bfi_cap, hbfi_cap; void *d_bfi; uint8_t *h_bfi;
    int pending;                /* work queued and not yet waited for */
    int failed;                 /* a queued call failed: reported by the next wait */
    long long launches;
    /* profiling: level 1 = device span of the calls (CUDA events, first kernel -> end of the last call before a wait);
     * level 2 = per-kernel events as well, which makes every call synchronous */
    cudaEvent_t *ev; int n_ev; double prof_ms[N_KERNELS + 3];
    cudaEvent_t span_ev; int span_open;
} Ctx;

static Ctx *g_ctx[LC3B_MAX_DEV];
static pthread_mutex_t g_lock = PTHREAD_MUTEX_INITIALIZER; /* context creation, handle registry, call ids */
static char g_err[256];
static __thread int tl_device = -1;
static int g_prof = 0;
static int g_chunk_dev = 5 << 15, g_chunk_host = 25 << 13, g_chunk_forced = 0; /* forced: set by the caller, taken literally */ /* channel-frames per chunk: input on the device | in host memory */
static unsigned g_call_id = 0;

static int cu_fail(cudaError_t e, const char *what)
{
    if (e == cudaSuccess) return 0;
    snprintf(g_err, sizeof(g_err), "%s: %s", what, cudaGetErrorString(e));
    return 1;
}
#define CU(x) do { if (cu_fail((x), #x)) return 1; } while (0)

/* ------------------------------------------------------------------ handle registry
 * The reference owns no memory, so callers re-initialise or free an encoder block without telling anybody.  The registry
 * maps a block's address to the state slots it was given: lc3_enc_init on an address that is still registered returns the
 * old slots to their pool first, so the normal reference usage pattern leaks nothing. */
typedef struct { uintptr_t key; int device, channels, slot[2]; } RegEnt;
static RegEnt *g_reg; static size_t g_reg_cap, g_reg_n;
static size_t reg_hash(uintptr_t k) { return (size_t)((k >> 4) * 0x9E3779B97F4A7C15ull); }
static RegEnt *reg_find(uintptr_t key)
{
    if (!g_reg_cap) return NULL;
    for (size_t i = reg_hash(key) & (g_reg_cap - 1);; i = (i + 1) & (g_reg_cap - 1)) {
        if (g_reg[i].key == key) return &g_reg[i];
        if (g_reg[i].key == 0) return NULL;
    }
}
static int reg_put(uintptr_t key, int device, int channels, const int *slot)
{
    if ((g_reg_n + 1) * 2 > g_reg_cap) {
        size_t nc = g_reg_cap ? g_reg_cap * 2 : 8192;
        RegEnt *n = (RegEnt *)calloc(nc, sizeof(RegEnt));
        if (!n) return 1;
        for (size_t i = 0; i < g_reg_cap; i++)
            if (g_reg[i].key > 1) {
                size_t j = reg_hash(g_reg[i].key) & (nc - 1);
                while (n[j].key) j = (j + 1) & (nc - 1);
                n[j] = g_reg[i];
            }
        free(g_reg); g_reg = n; g_reg_cap = nc; /* tombstones (key 1) are dropped */
    }
    size_t i = reg_hash(key) & (g_reg_cap - 1);
    while (g_reg[i].key > 1 && g_reg[i].key != key) i = (i + 1) & (g_reg_cap - 1);
    if (g_reg[i].key <= 1) g_reg_n++;
    g_reg[i].key = key; g_reg[i].device = device; g_reg[i].channels = channels; g_reg[i].slot[0] = slot[0]; g_reg[i].slot[1] = slot[1];
    return 0;
}
static void reg_del(RegEnt *e) { e->key = 1; } /* tombstone: probing continues past it */

/* ------------------------------------------------------------------ per-device context */
static size_t pool_bytes(int cap, int words, int elem) { return (size_t)cap * words * elem; }

static int ctx_wait(Ctx *c)
{ /* everything queued on this device is done (the device must be current) */
    int bad = 0;
    if (c->pending) {
        bad = cu_fail(cudaStreamSynchronize(c->join), "wait for the queued batches");
        if (!bad) bad = cu_fail(cudaStreamSynchronize(c->up), "wait for the uploads");
        c->pending = 0;
        if (!bad && g_prof && c->span_open) {
            float ms = 0;
            const cudaEvent_t end = c->call_end_ev[(c->call_no - 1) & 1];
            if (cudaEventElapsedTime(&ms, c->span_ev, end) == cudaSuccess) c->prof_ms[N_KERNELS] += ms;
        }
        c->span_open = 0;
    }
    if (c->failed) { /* a call stopped half-way through its queuing: nothing is joined, wait for the whole device */
        cudaDeviceSynchronize();
        c->failed = 0; c->pending = 0; c->span_open = 0;
        bad = 1;
    }
    return bad;
}

static int pool_grow(Ctx *c, int need)
{ /* tile-interleaved arrays grow by appending tiles, so the old contents are a prefix */
    int cap = c->cap ? c->cap : 4096;
    while (cap < need) cap *= 2;
    if (cap == c->cap) return 0;
    if (ctx_wait(c)) return 1; /* the old arrays may be in use */
    struct { void **p; int words, elem; } a[6] = {
        {(void **)&c->cfg, CFG_N, 4}, {(void **)&c->cfg_b, CFG_N, 4}, {(void **)&c->st32, ST32_N, 4}, {(void **)&c->hist32, LC3K_MAX_N, 4},
        {(void **)&c->st16, ST16_N, 2}, {(void **)&c->hist16, LC3K_MAX_N, 2}};
    for (int i = 0; i < 6; i++) {
        void *n = NULL;
        CU(cudaMalloc(&n, pool_bytes(cap, a[i].words, a[i].elem)));
        CU(cudaMemsetAsync(n, 0, pool_bytes(cap, a[i].words, a[i].elem), c->up));
        if (c->cap) CU(cudaMemcpyAsync(n, *a[i].p, pool_bytes(c->cap, a[i].words, a[i].elem), cudaMemcpyDeviceToDevice, c->up));
        CU(cudaStreamSynchronize(c->up));
        if (c->cap) CU(cudaFree(*a[i].p));
        *a[i].p = n;
    }
    int32_t *nh = NULL;
    CU(cudaMallocHost((void **)&nh, pool_bytes(cap, CFG_N, 4))); /* pinned: uploads from it are asynchronous */
    memset(nh, 0, pool_bytes(cap, CFG_N, 4));
    if (c->cfg_host) { memcpy(nh, c->cfg_host, pool_bytes(c->cap, CFG_N, 4)); cudaFreeHost(c->cfg_host); }
    c->cfg_host = nh;
    int *fl = (int *)realloc(c->free_list, (size_t)cap * sizeof(int));
    if (!fl) return 1;
    c->free_list = fl;
    c->cap = cap;
    return 0;
}

static int ctx_create(Ctx *c, int device)
{ /* the device is current */
    c->device = device;
    pthread_mutex_init(&c->lock, NULL);
    CU(cudaStreamCreateWithFlags(&c->up, cudaStreamNonBlocking));
    CU(cudaStreamCreateWithFlags(&c->join, cudaStreamNonBlocking));
    int prio_lo = 0, prio_hi = 0;
    CU(cudaDeviceGetStreamPriorityRange(&prio_lo, &prio_hi));
    for (int p = 0; p < N_PIPES; p++) {
        /* the per-stream kernels are short on parallelism (one warp or lane per stream) and sit on the critical path of
         * the next chunk: at higher priority their blocks are placed as soon as SM resources free up, so they run
         * underneath the previous chunk's frame-parallel kernels instead of queueing behind their grids */
        CU(cudaStreamCreateWithPriority(&c->pipe[p].par, cudaStreamNonBlocking, prio_lo));
        CU(cudaStreamCreateWithPriority(&c->pipe[p].ser, cudaStreamNonBlocking, prio_hi));
        for (int k = 0; k < N_KERNELS; k++) CU(cudaEventCreateWithFlags(&c->pipe[p].state_ev[k], cudaEventDisableTiming));
        for (int k = 0; k < N_KERNELS + 2; k++) CU(cudaEventCreateWithFlags(&c->pipe[p].hop_ev[k], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&c->pipe[p].done_ev, cudaEventDisableTiming));
    }
    CU(cudaEventCreateWithFlags(&c->call_ev, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&c->cfg_ev, cudaEventDisableTiming));
    CU(cudaEventCreateWithFlags(&c->tab_ev, cudaEventDisableTiming));
    CU(cudaEventCreate(&c->call_end_ev[0]));
    CU(cudaEventCreate(&c->call_end_ev[1]));
    CU(cudaEventCreate(&c->span_ev));
    CU(cudaStreamCreateWithFlags(&c->h2d, cudaStreamNonBlocking));
    CU(cudaStreamCreateWithFlags(&c->d2h, cudaStreamNonBlocking));
    CU(cudaEventCreateWithFlags(&c->d2h_ev, cudaEventDisableTiming));
    c->dirty_lo = 1 << 30; c->dirty_hi = -1;
    c->n_active = N_PIPES_MIN; c->next_pipe = 0; c->last_pipe = -1;
    return pool_grow(c, 4096);
}

/* context of `device` (-1: the device selected with lc3plus_set_device, else the thread's current device); created on
 * first use.  Leaves that device current.  NULL: no usable device (g_err says why). */
static Ctx *ctx_get(int device)
{
    if (device < 0) device = tl_device;
    if (device < 0) {
        int n = 0;
        if (cu_fail(cudaGetDeviceCount(&n), "cudaGetDeviceCount") || n <= 0) {
            if (!g_err[0] || n <= 0) snprintf(g_err, sizeof(g_err), "no CUDA device: liblc3plus_b200 has no CPU path");
            return NULL;
        }
        if (cu_fail(cudaGetDevice(&device), "cudaGetDevice")) return NULL;
    }
    if (device >= LC3B_MAX_DEV) { snprintf(g_err, sizeof(g_err), "device %d: at most %d devices", device, LC3B_MAX_DEV); return NULL; }
    if (cu_fail(cudaSetDevice(device), "cudaSetDevice")) return NULL;
    Ctx *c = g_ctx[device];
    if (c) return c;
    pthread_mutex_lock(&g_lock);
    c = g_ctx[device];
    if (!c) {
        c = (Ctx *)calloc(1, sizeof(Ctx));
        if (c && ctx_create(c, device)) { c = NULL; /* leaked on purpose: half-made CUDA objects */ }
        __sync_synchronize();
        g_ctx[device] = c;
    }
    pthread_mutex_unlock(&g_lock);
    return c;
}

static int slot_alloc(Ctx *c)
{
    int s;
    if (c->n_free > 0) s = c->free_list[--c->n_free];
    else {
        if (c->next_fresh >= c->cap && pool_grow(c, c->next_fresh + 1)) return -1;
        s = c->next_fresh++;
    }
    if (c->n_reset == c->cap_reset) {
        int nc = c->cap_reset ? c->cap_reset * 2 : 1024;
        int32_t *r = (int32_t *)realloc(c->reset_list, (size_t)nc * 4);
        if (!r) return -1;
        c->reset_list = r; c->cap_reset = nc;
    }
    c->reset_list[c->n_reset++] = s;
    return s;
}

static void cfg_to_shadow(Ctx *c, struct LC3_Enc *e)
{
    if (c->cfg_ev_used) { cudaEventSynchronize(c->cfg_ev); c->cfg_ev_used = 0; } /* the last upload has read the shadow (short: uploads queue on their own stream) */
    for (int ch = 0; ch < e->cfg.channels; ch++) {
        int slot = e->slot[ch], tile = slot >> 5, lane = slot & 31;
        int32_t *w = c->cfg_host + (size_t)tile * CFG_N * 32 + lane;
        if (e->cfg.attack_reset[ch]) {
            /* set_bitrate switched the attack detector off since the last batch: the reference clears the detector's state right
             * there (fx/setup_enc_lc3.c:266-275).  Here the lane's generation word moves on and the pitch pass, which owns that
             * state, clears it when the word differs from the one it stored -- no extra kernel, nothing to wait for. */
            e->cfg.attack_reset[ch] = 0;
            e->cfg.ch[ch][CFG_ATT_GEN]++;
        }
        for (int k = 0; k < CFG_N; k++) w[k * 32] = e->cfg.ch[ch][k];
        if (tile < c->dirty_lo) c->dirty_lo = tile;
        if (tile > c->dirty_hi) c->dirty_hi = tile;
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

static int is_device_ptr(const void *p)
{
    struct cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return 0; }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

static int upload_list(Ctx *c, const int32_t *list, int n)
{
    if (ensure(&c->d_list, &c->list_cap, (size_t)n * 4, 0)) return 1;
    CU(cudaMemcpyAsync(c->d_list, list, (size_t)n * 4, cudaMemcpyHostToDevice, c->up));
    CU(cudaStreamSynchronize(c->up)); /* the list is pageable and reused */
    return 0;
}

/* Power-on state for freshly initialised encoders (changes what queued kernels may read: the device is drained first; a steady
 * stream of calls has none), then the configuration words: the whole shadow goes into the copy of the pool that no queued
 * kernel reads (at most two calls are in flight and the one before the previous has finished), asynchronously, and this call's
 * kernels take that copy -- a bitrate switch between two asynchronous calls costs one small upload, not a drain. */
static int flush_pending(Ctx *c, int *uploaded)
{
    if (c->n_reset) {
        if (ctx_wait(c)) return 1;
        if (upload_list(c, c->reset_list, c->n_reset)) return 1;
        CU((cudaError_t)lc3k_launch_reset(c->st32, c->st16, c->hist16, c->hist32, (const int32_t *)c->d_list, c->n_reset, c->up));
        CU(cudaStreamSynchronize(c->up));
        c->launches++;
        c->n_reset = 0;
    }
    if (c->dirty_hi >= c->dirty_lo) {
        const size_t n = (size_t)((c->next_fresh + 31) / 32) * CFG_N * 32; /* every tile in use: the other copy is a call behind */
        c->cfg_cur ^= 1;
        CU(cudaMemcpyAsync(c->cfg_cur ? c->cfg_b : c->cfg, c->cfg_host, n * 4, cudaMemcpyHostToDevice, c->up));
        CU(cudaEventRecord(c->cfg_ev, c->up));
        c->cfg_ev_used = 1;
        c->dirty_lo = 1 << 30; c->dirty_hi = -1;
        *uploaded = 1;
    }
    return 0;
}

/* ------------------------------------------------------------------ batch engine */
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

typedef struct { const void *h_pcm; void *h_out; } ChunkCopy; /* host buffers to copy chunk by chunk (NULL: not needed) */
typedef struct { int uploads; /* something was queued on `up` that the chunks must wait for */ int d2h; unsigned pipes; } CallState;

/* one bucket: encoders that share (fs_idx, dms).  idx[] lists their positions in encs[]. */
static int run_bucket(Ctx *g, int key, struct LC3_Enc *const *encs, const int *idx, int n, int n_frames, const void *d_pcm, size_t stride,
                      int bits, uint8_t *d_out, size_t ostride, const int *num_bytes_in, const ChunkCopy *cp, CallState *cs)
{
    const LC3Cfg *c0 = &encs[idx[0]]->cfg;
    const int N = c0->N, sb = bits == 16 ? 2 : 4;
    size_t n_lanes = 0;
    for (int i = 0; i < n; i++) n_lanes += encs[idx[i]]->cfg.channels;
    /* lane tables: built in scratch, compared with what the device already holds for this bucket */
    const size_t tab_bytes = n_lanes * 24;
    if (g->tab_tmp_cap < tab_bytes) {
        unsigned char *t = (unsigned char *)realloc(g->tab_tmp, tab_bytes + tab_bytes / 2);
        if (!t) return 1;
        g->tab_tmp = t; g->tab_tmp_cap = tab_bytes + tab_bytes / 2;
    }
    int32_t *t_slot = (int32_t *)g->tab_tmp, *t_ostride = t_slot + n_lanes;
    int64_t *t_poff = (int64_t *)(t_ostride + n_lanes), *t_ooff = t_poff + n_lanes;
    int any_bfi = 0, max_bytes = 0, any_ep = 0;
    size_t l = 0;
    for (int i = 0; i < n; i++) {
        const struct LC3_Enc *e = encs[idx[i]];
        int nb = lc3cfg_num_bytes(&e->cfg), boff = 0;
        any_ep |= e->cfg.epmode;
        for (int ch = 0; ch < e->cfg.channels; ch++, l++) {
            if (e->cfg.ch[ch][CFG_TARGET_BYTES] > max_bytes) max_bytes = e->cfg.ch[ch][CFG_TARGET_BYTES];
            t_slot[l] = e->slot[ch];
            t_poff[l] = (int64_t)((size_t)idx[i] * stride / sb) + (int64_t)ch * n_frames * N;
            t_ooff[l] = (int64_t)((size_t)idx[i] * ostride) + boff;
            t_ostride[l] = nb;
            boff += lc3cfg_out_bytes(&e->cfg, ch);
        }
        if (num_bytes_in)
            for (int f = 0; f < n_frames && !any_bfi; f++) any_bfi = num_bytes_in[(size_t)idx[i] * n_frames + f] == -1;
    }
    Tab *tb = &g->tab[key];
    if (tb->n != n_lanes || memcmp(tb->h, g->tab_tmp, tab_bytes)) {
        if (g->tab_ev_used) { CU(cudaEventSynchronize(g->tab_ev)); g->tab_ev_used = 0; } /* short: uploads have their own stream */
        if (tb->cap < tab_bytes) {
            size_t z = 0;
            if (ctx_wait(g)) return 1; /* the copies being replaced may be in use */
            if (tb->h) { cudaFreeHost(tb->h); tb->h = NULL; }
            for (int k = 0; k < 2; k++) if (tb->d[k]) { cudaFree(tb->d[k]); tb->d[k] = NULL; }
            tb->cap = 0; tb->n = 0;
            if (ensure((void **)&tb->h, &z, tab_bytes, 1)) return 1;
            for (int k = 0; k < 2; k++) { z = 0; if (ensure((void **)&tb->d[k], &z, tab_bytes, 0)) return 1; }
            tb->cap = tab_bytes;
        }
        memcpy(tb->h, g->tab_tmp, tab_bytes);
        tb->n = n_lanes;
        tb->cur ^= 1;
        CU(cudaMemcpyAsync(tb->d[tb->cur], tb->h, tab_bytes, cudaMemcpyHostToDevice, g->up));
        CU(cudaEventRecord(g->tab_ev, g->up));
        g->tab_ev_used = 1;
        cs->uploads = 1;
    }

    LC3K_Batch b;
    memset(&b, 0, sizeof(b));
    lc3cfg_fill_batch(c0, &b);
    b.bits = bits;
    b.stage_full = !(bits == 16 && N >= 160);
    b.n_lanes = (int)n_lanes;
    b.n_tiles = (int)((n_lanes + 31) / 32);
    b.max_bytes = max_bytes;
    b.slot = (const int32_t *)tb->d[tb->cur]; b.out_stride = b.slot + n_lanes;
    b.pcm_off = (const int64_t *)(b.out_stride + n_lanes); b.out_off = b.pcm_off + n_lanes;
    b.cfg = g->cfg_cur ? g->cfg_b : g->cfg; b.st32 = g->st32; b.st16 = g->st16; b.hist16 = g->hist16; b.hist32 = g->hist32;
    b.pcm = d_pcm; b.out = d_out;

    if (any_bfi) { /* [n_frames][n_lanes] flags; rare (the caller marks lost frames): one call at a time */
        const size_t nbf = (size_t)n_frames * n_lanes;
        if (ctx_wait(g)) return 1;
        if (ensure((void **)&g->h_bfi, &g->hbfi_cap, nbf, 1) || ensure(&g->d_bfi, &g->bfi_cap, nbf, 0)) return 1;
        l = 0;
        for (int i = 0; i < n; i++)
            for (int ch = 0; ch < encs[idx[i]]->cfg.channels; ch++, l++)
                for (int f = 0; f < n_frames; f++) g->h_bfi[(size_t)f * n_lanes + l] = num_bytes_in[(size_t)idx[i] * n_frames + f] == -1;
        CU(cudaMemcpyAsync(g->d_bfi, g->h_bfi, nbf, cudaMemcpyHostToDevice, g->up));
        cs->uploads = 1;
        b.bfi = (const uint8_t *)g->d_bfi;
    }
    if (cs->uploads) CU(cudaEventRecord(g->call_ev, g->up));

    /* frames per chunk: about chunk_frames channel-frames each, balanced over the chunks */
    int fc = ((cp && cp->h_pcm) ? g_chunk_host : g_chunk_dev) / (b.n_tiles * 32);
    if (fc < 8 && !g_chunk_forced) fc = 8; /* very wide buckets: the per-stream kernels load and store their state once per chunk, a
                                              few frames per chunk would be mostly that (100 000 streams: 72 -> M frames/s) */
    if (fc < 1) fc = 1;
    if (fc > n_frames) fc = n_frames;
    const int ramp = cp && cp->h_pcm && n_frames > fc;
    int n_chunks = 0;
    for (int rem = n_frames; rem > 0; n_chunks++) rem -= chunk_size(n_chunks, rem, fc, ramp);
    const size_t ftl = (size_t)fc * b.n_tiles * 32;
    const size_t need[N_WORK] = {(size_t)b.n_tiles * 32 * (size_t)(fc + 1) * N * sb, ftl * PO_N * 4, ftl * N * 4, ftl * FO_N * 4,
                                 ftl * (N / 4) * 4, ftl * RO_N * 4, ftl * 128 * 2, ftl * N * 2, ftl * BEN_N * 4};
    for (int k = 0; k < (n_chunks < g->n_active ? n_chunks : g->n_active); k++) {
        const int p = (g->next_pipe + k) % g->n_active;
        for (int i = 0; i < N_WORK; i++)
            if (g->pipe[p].work_cap[i] < need[i]) {
                if (ctx_wait(g)) return 1; /* the buffer being replaced may be in use */
                if (ensure(&g->pipe[p].work[i], &g->pipe[p].work_cap[i], need[i], 0)) return 1;
            }
    }

    const int EPC = 2 * N_KERNELS; /* profiling events per chunk: one before and one after every kernel */
    const int kprof = g_prof >= 2;
    if (kprof && g->n_ev < n_chunks * EPC) {
        cudaEvent_t *ne = (cudaEvent_t *)realloc(g->ev, (size_t)n_chunks * EPC * sizeof(cudaEvent_t));
        if (!ne) return 1;
        g->ev = ne;
        for (; g->n_ev < n_chunks * EPC; g->n_ev++) CU(cudaEventCreate(&g->ev[g->n_ev]));
    }
    int (*const launch[N_KERNELS])(const LC3K_Batch *, void *) = {lc3k_launch_stage, lc3k_launch_fir, lc3k_launch_scalar, lc3k_launch_pitch,
                                                                  lc3k_launch_front, lc3k_launch_sns, lc3k_launch_tns, lc3k_launch_rate, lc3k_launch_quant, lc3k_launch_back};
    for (int f0 = 0, c = 0, cf; f0 < n_frames; f0 += cf, c++) {
        cf = chunk_size(c, n_frames - f0, fc, ramp);
        g->seq++;
        const int p = g->next_pipe % g->n_active, pprev = g->last_pipe;
        g->next_pipe = (p + 1) % g->n_active;
        g->last_pipe = p;
        const cudaStream_t par = g->pipe[p].par, ser = g->pipe[p].ser;
        void *const *w = g->pipe[p].work;
        /* the chunk that used these work buffers last is done with them (every kernel of this chunk follows its first one,
         * which runs on `ser`); the call's uploads are in place */
        if (g->pipe[p].used) CU(cudaStreamWaitEvent(ser, g->pipe[p].done_ev, 0));
        if (cs->uploads && c == 0) CU(cudaStreamWaitEvent(ser, g->call_ev, 0)); /* later chunks follow chunk 0's first kernel */
        b.frame0 = f0;
        b.n_frames = cf;
        b.pcm_t = w[0]; b.pout = (int32_t *)w[1]; b.spec = (int32_t *)w[2]; b.fout = (int32_t *)w[3];
        b.en4 = (int32_t *)w[4]; b.rout = (int32_t *)w[5]; b.r12 = (int16_t *)w[6]; b.xq = (int16_t *)w[7]; b.ben = (int32_t *)w[8];
        const int copy_in = cp && cp->h_pcm;
        if (copy_in) { /* this chunk's columns of every stream's PCM rows, one rectangle per channel */
            for (int ch = 0; ch < c0->channels; ch++) {
                const size_t off = ((size_t)ch * n_frames + f0) * N * sb;
                CU(cudaMemcpy2DAsync((char *)d_pcm + off, stride, (const char *)cp->h_pcm + off, stride, (size_t)b.n_frames * N * sb,
                                     (size_t)n, cudaMemcpyHostToDevice, g->h2d));
            }
            CU(cudaEventRecord(g->pipe[p].hop_ev[N_KERNELS], g->h2d));
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
        int first = 1;
        for (int o = 0; o < N_KERNELS; o++) {
            const int k = order[o];
            const cudaStream_t st = k_stateful[k] ? ser : par;
            if ((bits != 16 && (k == K_FIR || k == K_SCALAR)) /* fused into K_PITCH for 24-bit input */
                || (N < 160 && (k == K_SNS || k == K_TNS))) { /* fused into K_FRONT for the short frames */
                if (k_stateful[k]) { /* keep the chunk-to-chunk chain of this kernel's state events unbroken */
                    if (pprev >= 0 && pprev != p) CU(cudaStreamWaitEvent(st, g->pipe[pprev].state_ev[k], 0));
                    CU(cudaEventRecord(g->pipe[p].state_ev[k], st));
                }
                continue;
            }
            if (first) { /* the first kernel (K_STAGE, on `ser`) waits for the chunk's input copy */
                if (copy_in) CU(cudaStreamWaitEvent(st, g->pipe[p].hop_ev[N_KERNELS], 0));
                if (g_prof && !g->span_open) { CU(cudaEventRecord(g->span_ev, st)); g->span_open = 1; }
                first = 0;
            }
            const int *dep = split ? dep_split[k] : dep_chain[k];
            for (int j = 0; j < 2; j++)
                if (dep[j] >= 0 && ran[dep[j]] && (k_stateful[dep[j]] ? ser : par) != st) CU(cudaStreamWaitEvent(st, g->pipe[p].hop_ev[dep[j]], 0));
            /* the same kernel of the chunk before (of this or an earlier call) has left the per-stream state it hands over */
            if (k_stateful[k] && pprev >= 0 && pprev != p) CU(cudaStreamWaitEvent(st, g->pipe[pprev].state_ev[k], 0));
            if (kprof) CU(cudaEventRecord(g->ev[c * EPC + 2 * k], st));
            CU((cudaError_t)launch[k](&b, st));
            if (kprof) CU(cudaEventRecord(g->ev[c * EPC + 2 * k + 1], st));
            CU(cudaEventRecord(g->pipe[p].hop_ev[k], st));
            ran[k] = 1;
            if (k_stateful[k]) CU(cudaEventRecord(g->pipe[p].state_ev[k], st));
        }
        /* K_BACK ran on `par`; error protection, if any lane asks for it, codes the finished payloads in place */
        if (any_ep) { CU((cudaError_t)lc3k_launch_fec(&b, par)); g->launches++; }
        CU(cudaEventRecord(g->pipe[p].done_ev, par));
        g->pipe[p].used = 1;
        cs->pipes |= 1u << p;
        if (cp && cp->h_out) {
            const size_t nb = (size_t)lc3cfg_num_bytes(c0), off = (size_t)f0 * nb;
            CU(cudaStreamWaitEvent(g->d2h, g->pipe[p].done_ev, 0));
            CU(cudaMemcpy2DAsync((char *)cp->h_out + off, ostride, d_out + off, ostride, (size_t)b.n_frames * nb, (size_t)n,
                                 cudaMemcpyDeviceToHost, g->d2h));
            cs->d2h = 1;
        }
        g->launches += (bits == 16 ? 8 : 6) + (N >= 160 ? 2 : 0); /* frame lengths >= 160: the front end is three kernels (tile MDCT, SNS, TNS) */
    }
    if (kprof) { /* per-kernel time = elapsed on the kernel's own stream; kernels of chunks in flight together share the SMs */
        for (int p = 0; p < N_PIPES; p++) if (cs->pipes & (1u << p)) CU(cudaStreamSynchronize(g->pipe[p].par));
        float ms;
        for (int c = 0; c < n_chunks; c++)
            for (int k = 0; k < N_KERNELS; k++) {
                if ((bits != 16 && (k == K_FIR || k == K_SCALAR)) || (N < 160 && (k == K_SNS || k == K_TNS))) continue;
                CU(cudaEventElapsedTime(&ms, g->ev[c * EPC + 2 * k], g->ev[c * EPC + 2 * k + 1]));
                g->prof_ms[k] += ms;
            }
        g->prof_ms[N_KERNELS + 2] += n_chunks;
    }
    return 0;
}

/* the part of a batch that lives on one device: idx[] lists the positions in encs[] */
typedef struct {
    Ctx *g; struct LC3_Enc *const *encs; const int *idx; int n, n_frames; const void *pcm; size_t stride; int bits; void *out;
    size_t ostride; const int *num_bytes; int rc;
} DevJob;

static int run_on_device(DevJob *j)
{
    Ctx *g = j->g;
    struct LC3_Enc *const *encs = j->encs;
    const int *idx = j->idx;
    const int n = j->n, n_frames = j->n_frames, bits = j->bits, sb = bits == 16 ? 2 : 4;
    const size_t stride = j->stride, ostride = j->ostride;
    CU(cudaSetDevice(g->device));
    /* at most two calls in flight per device: call k reuses the staging buffers of call k - 2 */
    const int par2 = (int)(g->call_no & 1);
    if (g->call_end_used[par2]) CU(cudaEventSynchronize(g->call_end_ev[par2]));
    size_t max_in = 0, max_out = 0;
    int uniform = 1, one_bucket = 1, contiguous = 1;
    const struct LC3_Enc *e0 = encs[idx[0]];
    for (int i = 0; i < n; i++) {
        struct LC3_Enc *e = encs[idx[i]];
        if (e->cfg.dms_idx != e0->cfg.dms_idx || e->cfg.fs_idx != e0->cfg.fs_idx || e->cfg.channels != e0->cfg.channels) one_bucket = 0;
        if (idx[i] != idx[0] + i) contiguous = 0;
        if (e->dirty) cfg_to_shadow(g, e);
        size_t in_b = (size_t)e->cfg.channels * n_frames * e->cfg.N * sb, out_b = (size_t)n_frames * lc3cfg_num_bytes(&e->cfg);
        if (i && (in_b != max_in || out_b != max_out)) uniform = 0;
        if (in_b > max_in) max_in = in_b;
        if (out_b > max_out) max_out = out_b;
    }
    CallState cs = {0, 0, 0};
    {
        unsigned seen = 0;
        int nb = 0;
        for (int i = 0; i < n; i++) {
            const unsigned bit = 1u << (encs[idx[i]]->cfg.dms_idx * 5 + encs[idx[i]]->cfg.fs_idx);
            if (!(seen & bit)) { seen |= bit; nb++; }
        }
        g->n_active = nb < N_PIPES_MIN ? N_PIPES_MIN : (nb > N_PIPES ? N_PIPES : nb);
    }
    if (flush_pending(g, &cs.uploads)) return 1;

    const int pcm_dev = is_device_ptr(j->pcm), out_dev = is_device_ptr(j->out);
    /* this device's streams span rows idx[0] .. idx[n-1] of the caller's buffers; staging keeps the caller's row numbering
     * relative to the first row, so the lane tables are built from idx[i] - idx[0] */
    const size_t row0 = (size_t)idx[0], rows = (size_t)(idx[n - 1] - idx[0] + 1);
    const char *pcm = (const char *)j->pcm + row0 * stride;
    char *out = (char *)j->out + row0 * ostride;
    const void *d_pcm = pcm;
    uint8_t *d_out = (uint8_t *)out;
    /* equal-sized streams of one (sample rate, frame size) in consecutive rows: host buffers are copied chunk by chunk, so
     * the copies of one chunk overlap the kernels of its neighbours; otherwise whole-buffer copies up front / at the end */
    const int chunked = uniform && one_bucket && contiguous;
    ChunkCopy cp = {chunked && !pcm_dev ? pcm : NULL, chunked && !out_dev ? out : NULL};
    if (!pcm_dev) {
        if (g->in_cap[par2] < rows * stride + max_in) { if (ctx_wait(g)) return 1; }
        if (ensure(&g->d_in[par2], &g->in_cap[par2], rows * stride + max_in, 0)) return 1;
        d_pcm = g->d_in[par2];
        if (!chunked) {
            if (uniform && contiguous) {
                CU(cudaMemcpy2DAsync(g->d_in[par2], stride, pcm, stride, max_in, rows, cudaMemcpyHostToDevice, g->up));
            } else if (contiguous) {
                /* rows of unequal length (mixed sample rates / frame sizes): one rectangle as wide as the longest row (a row
                 * is `stride` bytes apart from the next, so reading past a shorter row's samples stays inside the buffer)
                 * for all rows but the last, which is copied exactly */
                const size_t wid = max_in < stride ? max_in : stride;
                const size_t last_b = (size_t)encs[idx[n - 1]]->cfg.channels * n_frames * encs[idx[n - 1]]->cfg.N * sb;
                if (rows > 1) CU(cudaMemcpy2DAsync(g->d_in[par2], stride, pcm, stride, wid, rows - 1, cudaMemcpyHostToDevice, g->up));
                CU(cudaMemcpyAsync((char *)g->d_in[par2] + (rows - 1) * stride, pcm + (rows - 1) * stride, last_b, cudaMemcpyHostToDevice, g->up));
            } else {
                for (int i = 0; i < n; i++) {
                    const size_t in_b = (size_t)encs[idx[i]]->cfg.channels * n_frames * encs[idx[i]]->cfg.N * sb, off = (idx[i] - row0) * stride;
                    CU(cudaMemcpyAsync((char *)g->d_in[par2] + off, pcm + off, in_b, cudaMemcpyHostToDevice, g->up));
                }
            }
            cs.uploads = 1;
        }
    }
    if (!out_dev) {
        if (g->out_cap[par2] < rows * ostride + max_out) { if (ctx_wait(g)) return 1; }
        if (ensure(&g->d_out[par2], &g->out_cap[par2], rows * ostride + max_out, 0)) return 1;
        d_out = (uint8_t *)g->d_out[par2];
    }

    /* bucket by (fs_idx, dms_idx); positions relative to this device's first row */
    int *order = (int *)malloc((size_t)n * sizeof(int));
    struct LC3_Enc **sub = (struct LC3_Enc **)malloc(rows * sizeof(*sub));
    int *nb_rel = j->num_bytes ? (int *)j->num_bytes + row0 * n_frames : NULL;
    int rc = !order || !sub;
    if (!rc) {
        memset(sub, 0, rows * sizeof(*sub));
        for (int i = 0; i < n; i++) sub[idx[i] - row0] = encs[idx[i]];
    }
    for (int key = 0; key < N_BUCKETS && !rc; key++) {
        int m = 0;
        for (int i = 0; i < n; i++)
            if (encs[idx[i]]->cfg.dms_idx * 5 + encs[idx[i]]->cfg.fs_idx == key) order[m++] = (int)(idx[i] - row0);
        if (m) rc = run_bucket(g, key, sub, order, m, n_frames, d_pcm, stride, bits, d_out, ostride, nb_rel, &cp, &cs);
    }
    free(order); free(sub);
    if (rc) { g->failed = 1; }
    /* the call's end: all pipes it used, its output copies */
    for (int p = 0; p < N_PIPES; p++) if (cs.pipes & (1u << p)) CU(cudaStreamWaitEvent(g->join, g->pipe[p].done_ev, 0));
    if (cs.d2h) { CU(cudaEventRecord(g->d2h_ev, g->d2h)); CU(cudaStreamWaitEvent(g->join, g->d2h_ev, 0)); }
    if (!rc && !out_dev && !chunked) {
        if (uniform && contiguous) {
            CU(cudaMemcpy2DAsync(out, ostride, d_out, ostride, max_out, rows, cudaMemcpyDeviceToHost, g->join));
        } else if (contiguous) {
            /* unequal rows: one rectangle as wide as the longest row for all rows but the last (the header says that the bytes
             * of a row past its own frames, up to the longest row of the batch, are overwritten), the last row exactly */
            const size_t wid = max_out < ostride ? max_out : ostride;
            const size_t last_b = (size_t)n_frames * lc3cfg_num_bytes(&encs[idx[n - 1]]->cfg);
            if (rows > 1) CU(cudaMemcpy2DAsync(out, ostride, d_out, ostride, wid, rows - 1, cudaMemcpyDeviceToHost, g->join));
            CU(cudaMemcpyAsync(out + (rows - 1) * ostride, d_out + (rows - 1) * ostride, last_b, cudaMemcpyDeviceToHost, g->join));
        } else {
            for (int i = 0; i < n; i++) {
                const size_t out_b = (size_t)n_frames * lc3cfg_num_bytes(&encs[idx[i]]->cfg), off = (idx[i] - row0) * ostride;
                CU(cudaMemcpyAsync(out + off, d_out + off, out_b, cudaMemcpyDeviceToHost, g->join));
            }
        }
    }
    CU(cudaEventRecord(g->call_end_ev[par2], g->join));
    g->call_end_used[par2] = 1;
    g->call_no++;
    g->pending = 1;
    if (g_prof) g->prof_ms[N_KERNELS + 1] += 1;
    if (g_prof >= 2 && ctx_wait(g)) return 1; /* per-kernel profiling: one call at a time */
    return rc;
}

static void *dev_thread(void *arg)
{
    DevJob *j = (DevJob *)arg;
    pthread_mutex_lock(&j->g->lock);
    j->rc = run_on_device(j);
    if (j->rc) j->g->failed = 1; /* whatever was queued before the failure is waited for (ctx_wait) before anything is reused */
    pthread_mutex_unlock(&j->g->lock);
    return NULL;
}

static int wait_device(Ctx *g)
{
    int bad = cu_fail(cudaSetDevice(g->device), "cudaSetDevice");
    pthread_mutex_lock(&g->lock);
    if (!bad) bad = ctx_wait(g);
    pthread_mutex_unlock(&g->lock);
    return bad;
}

static LC3_Error run_batch(struct LC3_Enc *const *encs, int n_streams, int n_frames, const void *pcm, size_t stride, int bits,
                           void *out, size_t ostride, int *num_bytes, int async)
{
    if (!encs || !pcm || !out) return LC3_NULL_ERROR;
    if (n_streams <= 0 || n_frames <= 0 || (bits != 16 && bits != 24)) return LC3_ERROR;
    const int sb = bits == 16 ? 2 : 4;
    int n_dev_used = 0, dev_of[LC3B_MAX_DEV], cnt[LC3B_MAX_DEV] = {0};
    pthread_mutex_lock(&g_lock);
    const unsigned id = ++g_call_id;
    pthread_mutex_unlock(&g_lock);
    for (int i = 0; i < n_streams; i++) {
        struct LC3_Enc *e = encs[i];
        if (!e) return LC3_NULL_ERROR;
        if (e->magic != LC3B_MAGIC) return LC3_ERROR;
        if (!e->cfg.br_set) return LC3_BITRATE_UNSET_ERROR;
        if (e->last_bits && e->last_bits != bits) {
            /* the reference carries the MDCT overlap across an lc3_enc16 <-> lc3_enc24 switch (fx/enc_lc3.c:35-83); here the two
             * input widths keep their history in different pools, so the switch is refused rather than coded differently */
            snprintf(g_err, sizeof(g_err), "encoder %d: bit depth changed from %d to %d (one encoder, one bit depth: re-initialise it)", i, e->last_bits, bits);
            return LC3_ERROR;
        }
        if ((unsigned)e->batch_mark == id) { snprintf(g_err, sizeof(g_err), "encoder listed twice in one batch"); return LC3_ERROR; }
        e->batch_mark = (int)id;
        const size_t in_b = (size_t)e->cfg.channels * n_frames * e->cfg.N * sb, out_b = (size_t)n_frames * lc3cfg_num_bytes(&e->cfg);
        if (n_streams > 1 && (stride < in_b || ostride < out_b)) { snprintf(g_err, sizeof(g_err), "stream stride too small"); return LC3_ERROR; }
        if (e->device < 0 || e->device >= LC3B_MAX_DEV || !g_ctx[e->device]) return LC3_ERROR;
        if (!cnt[e->device]++) dev_of[n_dev_used++] = e->device;
    }
    int cur = -1;
    cudaGetDevice(&cur);
    LC3_Error ret = LC3_ERROR;
    int *idx = (int *)malloc((size_t)n_streams * sizeof(int));
    DevJob jobs[LC3B_MAX_DEV];
    if (!idx) goto done;
    if (n_dev_used > 1 && (is_device_ptr(pcm) || is_device_ptr(out))) {
        snprintf(g_err, sizeof(g_err), "a batch whose encoders live on several devices needs host buffers (one call per device for device buffers)");
        goto done;
    }
    for (int d = 0, base = 0; d < n_dev_used; base += cnt[dev_of[d]], d++) {
        int m = 0;
        for (int i = 0; i < n_streams; i++) if (encs[i]->device == dev_of[d]) idx[base + m++] = i;
        DevJob j = {g_ctx[dev_of[d]], encs, idx + base, m, n_frames, pcm, stride, bits, out, ostride, num_bytes, 0};
        jobs[d] = j;
    }
    int bad = 0;
    if (n_dev_used == 1) { dev_thread(&jobs[0]); bad = jobs[0].rc; }
    else { /* one host thread per device queues that device's share */
        pthread_t th[LC3B_MAX_DEV];
        for (int d = 0; d < n_dev_used; d++) if (pthread_create(&th[d], NULL, dev_thread, &jobs[d])) { dev_thread(&jobs[d]); th[d] = 0; }
        for (int d = 0; d < n_dev_used; d++) { if (th[d]) pthread_join(th[d], NULL); bad |= jobs[d].rc; }
    }
    if (!async || bad)
        for (int d = 0; d < n_dev_used; d++) bad |= wait_device(jobs[d].g);
    if (bad) goto done;
    for (int i = 0; i < n_streams; i++) encs[i]->last_bits = bits;
    if (num_bytes)
        for (int i = 0; i < n_streams; i++) {
            int nb = lc3cfg_num_bytes(&encs[i]->cfg);
            for (int f = 0; f < n_frames; f++) num_bytes[(size_t)i * n_frames + f] = nb;
        }
    ret = LC3_OK;
done:
    free(idx);
    if (cur >= 0) cudaSetDevice(cur);
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

static void slots_free(int device, int channels, const int *slot)
{ /* g_lock held */
    Ctx *c = device >= 0 && device < LC3B_MAX_DEV ? g_ctx[device] : NULL;
    if (!c) return;
    pthread_mutex_lock(&c->lock);
    for (int ch = 0; ch < channels; ch++) c->free_list[c->n_free++] = slot[ch];
    pthread_mutex_unlock(&c->lock);
}

LC3_Error lc3_enc_init(LC3_Enc *encoder, int samplerate, int channels)
{ /* fx/lc3.c:124-131: NULL -> alignment -> samplerate -> channels */
    if (encoder == NULL) return LC3_NULL_ERROR;
    if ((uintptr_t)encoder % 4 != 0) return LC3_ALIGN_ERROR;
    if (!lc3_samplerate_supported(samplerate)) return LC3_SAMPLERATE_ERROR;
    if (!lc3_channels_supported(channels)) return LC3_CHANNELS_ERROR;
    /* a block that is initialised again (or was freed without lc3plus_enc_release and handed out again) gives its old
     * device state back first: the reference usage pattern leaks nothing */
    pthread_mutex_lock(&g_lock);
    RegEnt *old = reg_find((uintptr_t)encoder);
    if (old) { slots_free(old->device, old->channels, old->slot); reg_del(old); }
    pthread_mutex_unlock(&g_lock);
    memset(encoder, 0, sizeof(*encoder));
    int r = lc3cfg_init(&encoder->cfg, samplerate, channels);
    if (r) return (LC3_Error)r;
    int cur = -1;
    cudaGetDevice(&cur);
    Ctx *c = ctx_get(-1);
    int ok = c != NULL;
    if (ok) {
        pthread_mutex_lock(&c->lock);
        for (int ch = 0; ok && ch < channels; ch++) {
            encoder->slot[ch] = slot_alloc(c);
            if (encoder->slot[ch] < 0) ok = 0;
        }
        pthread_mutex_unlock(&c->lock);
        encoder->device = c->device;
    }
    if (ok) {
        pthread_mutex_lock(&g_lock);
        ok = !reg_put((uintptr_t)encoder, c->device, channels, encoder->slot);
        pthread_mutex_unlock(&g_lock);
    }
    if (cur >= 0) cudaSetDevice(cur);
    if (!ok) return LC3_ERROR; /* no device: there is no CPU path */
    encoder->magic = LC3B_MAGIC;
    return LC3_OK;
}

void lc3plus_enc_release(LC3_Enc *encoder)
{
    if (!encoder || encoder->magic != LC3B_MAGIC) return;
    pthread_mutex_lock(&g_lock);
    RegEnt *e = reg_find((uintptr_t)encoder);
    if (e) { slots_free(e->device, e->channels, e->slot); reg_del(e); }
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
{ /* fx/lc3.c:186-191.  The attack detector's state, which the reference clears right here when the new bitrate switches the
   * detector off (fx/setup_enc_lc3.c:266-275), is cleared on the device before the next batch (cfg_to_shadow). */
    if (encoder == NULL) return LC3_NULL_ERROR;
    if (encoder->magic != LC3B_MAGIC) return LC3_ERROR;
    int r = lc3cfg_set_bitrate(&encoder->cfg, bitrate);
    if (r == 0) encoder->dirty = 1;
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
static LC3_Error enc_one(LC3_Enc *encoder, void **in, int bits, void *out, int *num_bytes, void *scratch)
{ /* fx/lc3.c:246-257 */
    if (encoder == NULL || in == NULL || out == NULL || num_bytes == NULL || scratch == NULL) return LC3_NULL_ERROR;
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
    LC3_Error r = run_batch(list, 1, 1, planar, sizeof(planar), bits, frame, sizeof(frame), &nb, 0);
    if (r != LC3_OK) return r;
    memcpy(out, frame, (size_t)nb);
    *num_bytes = nb;
    return LC3_OK;
}
LC3_Error lc3_enc16(LC3_Enc *encoder, int16_t **input_samples, void *output_bytes, int *num_bytes, void *scratch)
{ /* fx/lc3.c:259-262 */
    return enc_one(encoder, (void **)input_samples, 16, output_bytes, num_bytes, scratch);
}
LC3_Error lc3_enc24(LC3_Enc *encoder, int32_t **input_samples, void *output_bytes, int *num_bytes, void *scratch)
{ /* fx/lc3.c:264-267 */
    return enc_one(encoder, (void **)input_samples, 24, output_bytes, num_bytes, scratch);
}

LC3_Error lc3plus_enc_batch(LC3_Enc *const *encoders, int n_streams, int n_frames, const void *pcm, size_t stream_stride_bytes,
                            int bitdepth, void *out, size_t out_stream_stride_bytes, int *num_bytes)
{
    return run_batch(encoders, n_streams, n_frames, pcm, stream_stride_bytes, bitdepth, out, out_stream_stride_bytes, num_bytes, 0);
}
LC3_Error lc3plus_enc_batch_async(LC3_Enc *const *encoders, int n_streams, int n_frames, const void *pcm, size_t stream_stride_bytes,
                                  int bitdepth, void *out, size_t out_stream_stride_bytes, int *num_bytes)
{
    return run_batch(encoders, n_streams, n_frames, pcm, stream_stride_bytes, bitdepth, out, out_stream_stride_bytes, num_bytes, 1);
}
LC3_Error lc3plus_sync(void)
{
    int cur = -1, bad = 0;
    cudaGetDevice(&cur);
    for (int d = 0; d < LC3B_MAX_DEV; d++) if (g_ctx[d]) bad |= wait_device(g_ctx[d]);
    if (cur >= 0) cudaSetDevice(cur);
    return bad ? LC3_ERROR : LC3_OK;
}
LC3_Error lc3plus_set_device(int device)
{
    if (device < 0) { tl_device = -1; return LC3_OK; }
    int n = 0;
    if (cu_fail(cudaGetDeviceCount(&n), "cudaGetDeviceCount") || device >= n || device >= LC3B_MAX_DEV) {
        snprintf(g_err, sizeof(g_err), "lc3plus_set_device(%d): no such device", device);
        return LC3_ERROR;
    }
    tl_device = device;
    return LC3_OK;
}
int lc3plus_enc_get_device(const LC3_Enc *encoder) { return enc_ok(encoder) ? encoder->device : -1; }
/* for lc3_dec_api.c: the device a new handle goes to (lc3plus_set_device, else the thread's current device); -1: none */
int lc3b_pick_device(void)
{
    int device = tl_device;
    if (device < 0) {
        int n = 0;
        if (cu_fail(cudaGetDeviceCount(&n), "cudaGetDeviceCount") || n <= 0) {
            if (!g_err[0] || n <= 0) snprintf(g_err, sizeof(g_err), "no CUDA device: liblc3plus_b200 has no CPU path");
            return -1;
        }
        if (cu_fail(cudaGetDevice(&device), "cudaGetDevice")) return -1;
    }
    return device;
}

long long lc3plus_kernel_launches(void)
{
    long long n = 0;
    for (int d = 0; d < LC3B_MAX_DEV; d++) if (g_ctx[d]) n += g_ctx[d]->launches;
    return n;
}
void lc3plus_profile_enable(int level)
{ /* 0 off, 1 device span of the calls, 2 per-kernel events as well (every call then waits for the device) */
    lc3plus_sync();
    g_prof = level;
    for (int d = 0; d < LC3B_MAX_DEV; d++) if (g_ctx[d]) memset(g_ctx[d]->prof_ms, 0, sizeof(g_ctx[d]->prof_ms));
}
void lc3plus_profile_read(double *ms)
{ /* summed over the devices in use (a benchmark process drives one) */
    lc3plus_sync();
    memset(ms, 0, sizeof(double) * (N_KERNELS + 3));
    for (int d = 0; d < LC3B_MAX_DEV; d++)
        if (g_ctx[d]) for (int k = 0; k < N_KERNELS + 3; k++) ms[k] += g_ctx[d]->prof_ms[k];
}
void lc3plus_profile_read_device(int device, double *ms)
{
    lc3plus_sync();
    memset(ms, 0, sizeof(double) * (N_KERNELS + 3));
    if (device >= 0 && device < LC3B_MAX_DEV && g_ctx[device]) memcpy(ms, g_ctx[device]->prof_ms, sizeof(double) * (N_KERNELS + 3));
}
LC3_Error lc3plus_enc_set_bitrate_many(LC3_Enc *const *encoders, int n, const int *bitrates)
{ /* lc3_enc_set_bitrate on each encoder (fx/codec_exe.c:340-349 does this per frame from its switching file) */
    if (!encoders || !bitrates) return LC3_NULL_ERROR;
    LC3_Error first = LC3_OK;
    for (int i = 0; i < n; i++) {
        const LC3_Error r = lc3_enc_set_bitrate(encoders[i], bitrates[i]);
        if (r != LC3_OK && first == LC3_OK) first = r;
    }
    return first;
}
const char *lc3plus_last_error(void) { return g_err; }
void lc3b_set_error(const char *msg) { snprintf(g_err, sizeof(g_err), "%s", msg); } /* for lc3_k_noalloc.cu */
void lc3plus_set_chunk_frames(int channel_frames)
{ /* >= 32: both chunk sizes; 0: back to the defaults */
    if (channel_frames >= 32) { g_chunk_dev = g_chunk_host = channel_frames; g_chunk_forced = 1; }
    else if (channel_frames == 0) { g_chunk_dev = 5 << 15; g_chunk_host = 25 << 13; g_chunk_forced = 0; }
    { const char *e = getenv("LC3PLUS_B200_HOST_CHUNK"); if (channel_frames == 0 && e && atoi(e) >= 32) g_chunk_host = atoi(e); } /* tuning aid */
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
