/* Host C layer of the DECODER half of liblc3plus_b200: the reference decoder API (fx/lc3.h:334-466,
 * fx/lc3.c:270-420) for error-free streams and the batched entry point lc3plus_dec_batch on top of the CUDA kernels of
 * lc3_k_dec.cu.  Plain C; there is no CPU decode path in this library.
 *
 * Scope: good frames only.  Packet-loss and partial concealment (fx/plc_*.c, fx/pc_*.c) and the channel decoder
 * (fx/al_fec.c fec_decoder) are not built: a payload the reference would conceal is synthesised as a zero spectrum, flagged
 * per frame, and the call returns LC3_DECODE_ERROR; lc3_dec_set_ep_enabled(1) is refused.
 *
 * One context per device, as the encoder's (lc3_api.c), but synchronous: a call returns when its PCM is in place.  Inside a
 * call with host buffers the payload uploads run ahead of the kernels and the PCM downloads behind them, chunk by chunk. */
#include <cuda_runtime_api.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "../../include/lc3plus_b200.h"
#include "lc3_dec_config.h"

#define LC3D_MAGIC 0x4c433364u
#define LC3D_MAX_DEV 16
enum { W_XQ = 0, W_SPEC, W_DOUT, W_YROW, W_WORK, N_DWORK };

struct LC3_Dec { /* lives in the caller's block; device state is referenced by slot */
    uint32_t magic;
    int device;
    int slot[2];
    int batch_mark;
    LC3DCfg cfg;
};

typedef struct {
    int device;
    pthread_mutex_t lock;
    cudaStream_t s, cp, dl;        /* kernels | payload uploads of the chunks ahead | PCM downloads of the chunks behind */
    cudaEvent_t *in_ev, *k_ev; int n_ev; /* per chunk of a call: its payloads are on the device | its PCM is ready */
    int cap;                       /* slots, multiple of 32 */
    int32_t *st32; int16_t *st16;  /* device state pool, tile-interleaved (lc3_dec_layout.h) */
    int *free_list, n_free, cap_free, next_fresh;
    int32_t *reset_list; int n_reset, cap_reset;
    void *d_list; size_t list_cap;
    size_t work_cap[N_DWORK]; void *work[N_DWORK];   /* per-chunk work buffers (the chunks of a call run one after another) */
    size_t in_cap, out_cap, bfi_cap; void *d_in, *d_out; uint8_t *d_bfi, *h_bfi; size_t hbfi_cap; /* staging that mirrors host buffers */
    unsigned char *tab_h, *tab_d; size_t tab_cap;
    long long launches;
    double prof_ms[4]; int prof; cudaEvent_t pev[5];
} DCtx;

static DCtx *g_dctx[LC3D_MAX_DEV];
static pthread_mutex_t g_dlock = PTHREAD_MUTEX_INITIALIZER;
static unsigned g_dcall = 0;
static int g_dchunk = 5 << 15; /* channel-frames per chunk */

void lc3b_set_error(const char *msg);   /* lc3_api.c */
int lc3b_pick_device(void);             /* lc3_api.c: the device selected with lc3plus_set_device, else the thread's current one; -1: none */

static int dfail(cudaError_t e, const char *what)
{
    if (e == cudaSuccess) return 0;
    char m[200];
    snprintf(m, sizeof(m), "%s: %s", what, cudaGetErrorString(e));
    lc3b_set_error(m);
    return 1;
}
#define CU(x) do { if (dfail((x), #x)) return 1; } while (0)

/* ---- handle registry (as lc3_api.c: the caller re-initialises or frees a decoder block without telling anybody) */
typedef struct { uintptr_t key; int device, channels, slot[2]; } DReg;
static DReg *g_reg; static size_t g_reg_cap, g_reg_n;
static size_t reg_hash(uintptr_t k) { return (size_t)((k >> 4) * 0x9E3779B97F4A7C15ull); }
static DReg *reg_find(uintptr_t key)
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
        size_t nc = g_reg_cap ? g_reg_cap * 2 : 4096;
        DReg *n = (DReg *)calloc(nc, sizeof(DReg));
        if (!n) return 1;
        for (size_t i = 0; i < g_reg_cap; i++)
            if (g_reg[i].key > 1) {
                size_t j = reg_hash(g_reg[i].key) & (nc - 1);
                while (n[j].key) j = (j + 1) & (nc - 1);
                n[j] = g_reg[i];
            }
        free(g_reg); g_reg = n; g_reg_cap = nc;
        g_reg_n = 0;
        for (size_t i = 0; i < nc; i++) g_reg_n += n[i].key > 1;
    }
    size_t i = reg_hash(key) & (g_reg_cap - 1);
    while (g_reg[i].key > 1 && g_reg[i].key != key) i = (i + 1) & (g_reg_cap - 1);
    if (g_reg[i].key <= 1) g_reg_n++;
    g_reg[i].key = key; g_reg[i].device = device; g_reg[i].channels = channels; g_reg[i].slot[0] = slot[0]; g_reg[i].slot[1] = slot[1];
    return 0;
}

/* ---- per-device context */
static int pool_grow(DCtx *c, int want)
{
    int cap = c->cap ? c->cap : 1024;
    while (cap < want) cap *= 2;
    if (cap == c->cap) return 0;
    int32_t *n32; int16_t *n16;
    CU(cudaMalloc((void **)&n32, (size_t)cap * DST32_N * 4));
    CU(cudaMalloc((void **)&n16, (size_t)cap * DST16_N * 2));
    if (c->cap) { /* tiles are contiguous: the old pool is a prefix of the new one */
        CU(cudaStreamSynchronize(c->s));
        CU(cudaMemcpy(n32, c->st32, (size_t)c->cap * DST32_N * 4, cudaMemcpyDeviceToDevice));
        CU(cudaMemcpy(n16, c->st16, (size_t)c->cap * DST16_N * 2, cudaMemcpyDeviceToDevice));
        cudaFree(c->st32); cudaFree(c->st16);
    }
    c->st32 = n32; c->st16 = n16; c->cap = cap;
    return 0;
}
static DCtx *dctx_get(int device)
{
    if (device < 0) device = lc3b_pick_device();
    if (device < 0) return NULL;
    if (device >= LC3D_MAX_DEV) { lc3b_set_error("decoder: device index beyond the supported 16"); return NULL; }
    if (dfail(cudaSetDevice(device), "cudaSetDevice")) return NULL;
    if (g_dctx[device]) return g_dctx[device];
    DCtx *c = (DCtx *)calloc(1, sizeof(DCtx));
    if (!c) return NULL;
    c->device = device;
    pthread_mutex_init(&c->lock, NULL);
    if (dfail(cudaStreamCreateWithFlags(&c->s, cudaStreamNonBlocking), "cudaStreamCreate") ||
        dfail(cudaStreamCreateWithFlags(&c->cp, cudaStreamNonBlocking), "cudaStreamCreate") ||
        dfail(cudaStreamCreateWithFlags(&c->dl, cudaStreamNonBlocking), "cudaStreamCreate")) { free(c); return NULL; }
    for (int i = 0; i < 5; i++) if (dfail(cudaEventCreate(&c->pev[i]), "cudaEventCreate")) { free(c); return NULL; }
    if (pool_grow(c, 1024)) { free(c); return NULL; }
    g_dctx[device] = c;
    return c;
}
static int slot_alloc(DCtx *c)
{
    if (c->n_free > 0) return c->free_list[--c->n_free];
    if (c->next_fresh >= c->cap && pool_grow(c, c->cap * 2)) return -1;
    return c->next_fresh++;
}
static int slot_free(DCtx *c, int slot)
{
    if (c->n_free >= c->cap_free) {
        int nc = c->cap_free ? c->cap_free * 2 : 1024;
        int *n = (int *)realloc(c->free_list, (size_t)nc * sizeof(int));
        if (!n) return 1;
        c->free_list = n; c->cap_free = nc;
    }
    c->free_list[c->n_free++] = slot;
    return 0;
}
static int mark_reset(DCtx *c, int slot)
{
    if (c->n_reset >= c->cap_reset) {
        int nc = c->cap_reset ? c->cap_reset * 2 : 1024;
        int32_t *n = (int32_t *)realloc(c->reset_list, (size_t)nc * 4);
        if (!n) return 1;
        c->reset_list = n; c->cap_reset = nc;
    }
    c->reset_list[c->n_reset++] = slot;
    return 0;
}
static int ensure(void **p, size_t *cap, size_t need)
{
    if (*cap >= need) return 0;
    if (*p) cudaFree(*p);
    *p = NULL; *cap = 0;
    size_t n = need + need / 4;
    CU(cudaMalloc(p, n));
    *cap = n;
    return 0;
}
static int ensure_events(DCtx *c, int n)
{
    if (n <= c->n_ev) return 0;
    cudaEvent_t *a = (cudaEvent_t *)realloc(c->in_ev, (size_t)n * sizeof(cudaEvent_t));
    if (a) c->in_ev = a;
    cudaEvent_t *k = (cudaEvent_t *)realloc(c->k_ev, (size_t)n * sizeof(cudaEvent_t));
    if (k) c->k_ev = k;
    if (!a || !k) return 1;
    for (; c->n_ev < n; c->n_ev++) {
        CU(cudaEventCreateWithFlags(&c->in_ev[c->n_ev], cudaEventDisableTiming));
        CU(cudaEventCreateWithFlags(&c->k_ev[c->n_ev], cudaEventDisableTiming));
    }
    return 0;
}
static int flush_resets(DCtx *c)
{
    if (!c->n_reset) return 0;
    if (ensure(&c->d_list, &c->list_cap, (size_t)c->n_reset * 4)) return 1;
    CU(cudaMemcpyAsync(c->d_list, c->reset_list, (size_t)c->n_reset * 4, cudaMemcpyHostToDevice, c->s));
    int e = lc3d_launch_reset(c->st32, c->st16, (const int32_t *)c->d_list, c->n_reset, c->s);
    if (e) return dfail((cudaError_t)e, "decoder state reset");
    CU(cudaStreamSynchronize(c->s)); /* reset_list is pageable */
    c->launches++;
    c->n_reset = 0;
    return 0;
}

static int dec_ok(const LC3_Dec *d) { return d && d->magic == LC3D_MAGIC; }

/* ------------------------------------------------------------------ reference API (fx/lc3.h:334-466) */
int lc3_dec_get_size(int samplerate, int channels, LC3_PlcMode plc_mode)
{
    if (!lc3_samplerate_supported(samplerate) || !lc3_channels_supported(channels)) return 0;
    if (plc_mode != LC3_PLC_STANDARD && plc_mode != LC3_PLC_ADVANCED) return 0;
    return (int)sizeof(struct LC3_Dec);
}
LC3_Error lc3_dec_init(LC3_Dec *dec, int samplerate, int channels, LC3_PlcMode plc_mode)
{
    if (!dec) return LC3_NULL_ERROR;
    if (!lc3_samplerate_supported(samplerate)) return LC3_SAMPLERATE_ERROR;
    if (!lc3_channels_supported(channels)) return LC3_CHANNELS_ERROR;
    if (plc_mode != LC3_PLC_STANDARD && plc_mode != LC3_PLC_ADVANCED) return LC3_PLCMODE_ERROR;
    LC3DCfg cfg;
    int r = lc3dcfg_init(&cfg, samplerate, channels);
    if (r) return (LC3_Error)r;
    pthread_mutex_lock(&g_dlock);
    DCtx *c = dctx_get(-1);
    if (!c) { pthread_mutex_unlock(&g_dlock); return LC3_ERROR; }
    DReg *old = reg_find((uintptr_t)dec);
    if (old) { /* the block is being re-initialised: its old slots go back to their pool */
        DCtx *oc = g_dctx[old->device];
        pthread_mutex_lock(&oc->lock);
        for (int ch = 0; ch < old->channels; ch++) slot_free(oc, old->slot[ch]);
        pthread_mutex_unlock(&oc->lock);
        old->key = 1; /* tombstone */
    }
    memset(dec, 0, sizeof(*dec));
    dec->cfg = cfg;
    dec->device = c->device;
    dec->slot[0] = dec->slot[1] = -1;
    int bad = 0;
    pthread_mutex_lock(&c->lock); /* lock order: registry, then context */
    for (int ch = 0; ch < channels && !bad; ch++) {
        dec->slot[ch] = slot_alloc(c);
        bad = dec->slot[ch] < 0 || mark_reset(c, dec->slot[ch]);
    }
    pthread_mutex_unlock(&c->lock);
    if (!bad) bad = reg_put((uintptr_t)dec, c->device, channels, dec->slot);
    pthread_mutex_unlock(&g_dlock);
    if (bad) return LC3_ERROR;
    dec->magic = LC3D_MAGIC;
    return LC3_OK;
}
LC3_Error lc3_dec_free(LC3_Dec *dec)
{ /* NEW: returns the decoder's device state to the pool (optional: re-initialising or abandoning the block leaks nothing) */
    if (!dec_ok(dec)) return LC3_NULL_ERROR;
    pthread_mutex_lock(&g_dlock);
    DReg *e = reg_find((uintptr_t)dec);
    if (e) {
        DCtx *oc = g_dctx[e->device];
        pthread_mutex_lock(&oc->lock);
        for (int ch = 0; ch < e->channels; ch++) slot_free(oc, e->slot[ch]);
        pthread_mutex_unlock(&oc->lock);
        e->key = 1;
    }
    dec->magic = 0;
    pthread_mutex_unlock(&g_dlock);
    return LC3_OK;
}
int lc3_dec_get_scratch_size(const LC3_Dec *dec) { return dec_ok(dec) ? 16 : 0; } /* the scratch argument is accepted and not used */
int lc3_dec_get_output_samples(const LC3_Dec *dec) { return dec_ok(dec) ? dec->cfg.N : 0; }
int lc3_dec_get_delay(const LC3_Dec *dec) { return dec_ok(dec) ? lc3dcfg_delay(&dec->cfg) : 0; }
LC3_Error lc3_dec_set_frame_dms(LC3_Dec *dec, int frame_dms)
{
    if (!dec_ok(dec)) return LC3_NULL_ERROR;
    return (LC3_Error)lc3dcfg_set_frame_dms(&dec->cfg, frame_dms);
}
LC3_Error lc3_dec_set_ep_enabled(LC3_Dec *dec, int ep_enabled)
{
    if (!dec_ok(dec)) return LC3_NULL_ERROR;
    if (ep_enabled) { lc3b_set_error("lc3_dec_set_ep_enabled: the channel decoder is not part of this library"); return LC3_EPMODE_ERROR; }
    return LC3_OK;
}
LC3_EpModeRequest lc3_dec_get_ep_mode_request(const LC3_Dec *dec) { (void)dec; return LC3_EPMR_ZERO; }
int lc3_dec_get_error_report(const LC3_Dec *dec) { (void)dec; return 0; }
int lc3_dec_get_epok_flags(const LC3_Dec *dec) { (void)dec; return 0; }
int lc3plus_dec_get_device(const LC3_Dec *dec) { return dec_ok(dec) ? dec->device : -1; }

/* ------------------------------------------------------------------ batch */
static int is_dev_ptr(const void *p)
{
    struct cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return 0; }
    return a.type == cudaMemoryTypeDevice || a.type == cudaMemoryTypeManaged;
}

/* lane tables of a call, one pinned block mirrored to the device:
 * cfg[n_tiles][DCFG_N][32] i32 | slot[n] i32 | in_stride[n] i32 | in_off[n] i64 | pcm_off[n] i64 */
typedef struct { size_t cfg, slot, stride, ioff, poff, total; } TabLay;
static TabLay tab_layout(int n_lanes, int n_tiles)
{
    TabLay t;
    size_t o = 0;
    t.cfg = o; o += (size_t)n_tiles * DCFG_N * 32 * 4;
    t.slot = o; o += ((size_t)n_lanes * 4 + 7) & ~(size_t)7;
    t.stride = o; o += ((size_t)n_lanes * 4 + 7) & ~(size_t)7;
    t.ioff = o; o += (size_t)n_lanes * 8;
    t.poff = o; o += (size_t)n_lanes * 8;
    t.total = o;
    return t;
}

/* One bucket: n_streams decoders that share sample rate, frame duration and device.  rows: the row of each of them in the
 * caller's buffers and arrays (NULL: decoder i is row i); n_rows: rows of the whole call. */
static LC3_Error dec_bucket(LC3_Dec *const *decs, const int *rows, int n_streams, int n_rows, int n_frames, const void *in, size_t in_stream_stride,
                            const int *num_bytes, void *pcm, size_t pcm_stream_stride, int bitdepth, uint8_t *bfi_out)
{
#define ROW(i) (rows ? rows[i] : (i))
    const LC3DCfg *g = &decs[0]->cfg;
    const int dev = decs[0]->device, N = g->N, sb = bitdepth == 16 ? 2 : 4;
    int n_lanes = 0;
    for (int i = 0; i < n_streams; i++) n_lanes += decs[i]->cfg.channels;
    pthread_mutex_lock(&g_dlock);
    DCtx *c = dctx_get(dev);
    pthread_mutex_unlock(&g_dlock);
    if (!c) return LC3_ERROR;
    LC3_Error ret = LC3_ERROR;
    pthread_mutex_lock(&c->lock);
    const int n_tiles = (n_lanes + 31) / 32;
    const TabLay tl = tab_layout(n_lanes, n_tiles);
    if (c->tab_cap < tl.total) {
        if (c->tab_h) cudaFreeHost(c->tab_h);
        if (c->tab_d) cudaFree(c->tab_d);
        c->tab_h = NULL; c->tab_d = NULL; c->tab_cap = 0;
        const size_t n = tl.total + tl.total / 2;
        if (dfail(cudaMallocHost((void **)&c->tab_h, n), "cudaMallocHost") || dfail(cudaMalloc((void **)&c->tab_d, n), "cudaMalloc")) goto out;
        c->tab_cap = n;
    }
    {
        int32_t *cfgw = (int32_t *)(c->tab_h + tl.cfg), *slot = (int32_t *)(c->tab_h + tl.slot), *stride = (int32_t *)(c->tab_h + tl.stride);
        int64_t *ioff = (int64_t *)(c->tab_h + tl.ioff), *poff = (int64_t *)(c->tab_h + tl.poff);
        memset(cfgw, 0, (size_t)n_tiles * DCFG_N * 32 * 4);
        int li = 0, max_bytes = 0;
        for (int i = 0; i < n_streams; i++) {
            const LC3DCfg *dc = &decs[i]->cfg;
            const int nb = num_bytes[ROW(i)];
            if (nb < 0 || (n_rows > 1 && (size_t)nb * n_frames > in_stream_stride)) { lc3b_set_error("lc3plus_dec_batch: num_bytes does not fit the input stride"); goto out; }
            int off = 0;
            for (int ch = 0; ch < dc->channels; ch++, li++) {
                const int cb = nb / dc->channels + (ch < nb % dc->channels); /* fx/dec_lc3.c:455 */
                int32_t w[DCFG_N] = {0, 0, 0, -1};
                if (nb > 0) {
                    const int r = lc3dcfg_lane_words(dc, cb, w);
                    if (r) { ret = (LC3_Error)r; goto out; }
                }
                for (int k = 0; k < DCFG_N; k++) cfgw[((size_t)(li >> 5) * DCFG_N + k) * 32 + (li & 31)] = w[k];
                slot[li] = decs[i]->slot[ch];
                stride[li] = nb;
                ioff[li] = (int64_t)ROW(i) * (int64_t)in_stream_stride + off;
                poff[li] = ((int64_t)ROW(i) * (int64_t)pcm_stream_stride) / sb + (int64_t)ch * n_frames * N;
                off += cb;
                if (cb > max_bytes) max_bytes = cb;
            }
            if (pcm_stream_stride % sb || (n_rows > 1 && pcm_stream_stride < (size_t)dc->channels * n_frames * N * sb)) {
                lc3b_set_error("lc3plus_dec_batch: pcm stride too small or misaligned"); goto out;
            }
        }
        if (flush_resets(c)) goto out;
        if (dfail(cudaMemcpyAsync(c->tab_d, c->tab_h, tl.total, cudaMemcpyHostToDevice, c->s), "table upload")) goto out;

        const int in_dev = is_dev_ptr(in), out_dev = is_dev_ptr(pcm);
        /* staging mirrors the caller's buffers: sized for all rows of the call (a row's own length is at most its stride, or
         * what the last row needs) */
        size_t in_bytes = 0, out_bytes = 0;
        for (int i = 0; i < n_streams; i++) {
            const size_t ie = (size_t)ROW(i) * in_stream_stride + (size_t)num_bytes[ROW(i)] * n_frames;
            const size_t oe = (size_t)ROW(i) * pcm_stream_stride + (size_t)decs[i]->cfg.channels * n_frames * N * sb;
            if (ie > in_bytes) in_bytes = ie;
            if (oe > out_bytes) out_bytes = oe;
        }
        const uint8_t *d_in = (const uint8_t *)in;
        void *d_pcm = pcm;
        /* host buffers, every stream with the same byte and channel count: the chunks' payloads and PCM move as rectangles
         * (rows = streams) on their own streams, uploads ahead of the kernels and downloads behind them */
        int rect = (!in_dev || !out_dev) && !rows;
        for (int i = 1; i < n_streams && rect; i++)
            rect = num_bytes[i] == num_bytes[0] && decs[i]->cfg.channels == decs[0]->cfg.channels;
        if (!in_dev) {
            if (ensure(&c->d_in, &c->in_cap, in_bytes)) goto out;
            if (!rect) /* row by row: the rows of other buckets and devices are not this bucket's business */
                for (int i = 0; i < n_streams; i++) {
                    const size_t o = (size_t)ROW(i) * in_stream_stride, len = (size_t)num_bytes[ROW(i)] * n_frames;
                    if (len && dfail(cudaMemcpyAsync((uint8_t *)c->d_in + o, (const uint8_t *)in + o, len, cudaMemcpyHostToDevice, c->s), "payload upload")) goto out;
                }
            d_in = (const uint8_t *)c->d_in;
        }
        if (!out_dev) {
            if (ensure(&c->d_out, &c->out_cap, out_bytes)) goto out;
            d_pcm = c->d_out;
        }
        const size_t bfi_bytes = (size_t)n_frames * n_lanes;
        if (ensure((void **)&c->d_bfi, &c->bfi_cap, bfi_bytes)) goto out;
        if (c->hbfi_cap < bfi_bytes) {
            if (c->h_bfi) cudaFreeHost(c->h_bfi);
            c->h_bfi = NULL; c->hbfi_cap = 0;
            if (dfail(cudaMallocHost((void **)&c->h_bfi, bfi_bytes + bfi_bytes / 2), "cudaMallocHost")) goto out;
            c->hbfi_cap = bfi_bytes + bfi_bytes / 2;
        }

        LC3D_Batch b;
        memset(&b, 0, sizeof(b));
        lc3dcfg_fill_batch(g, &b);
        b.bits = bitdepth; b.n_lanes = n_lanes; b.n_tiles = n_tiles; b.max_bytes = max_bytes;
        b.st32 = c->st32; b.st16 = c->st16;
        b.cfg = (const int32_t *)(c->tab_d + tl.cfg); b.slot = (const int32_t *)(c->tab_d + tl.slot);
        b.in_stride = (const int32_t *)(c->tab_d + tl.stride); b.in_off = (const int64_t *)(c->tab_d + tl.ioff);
        b.pcm_off = (const int64_t *)(c->tab_d + tl.poff);
        b.in = d_in; b.pcm = d_pcm; b.bfi = c->d_bfi;
        int chunk = g_dchunk / (n_tiles * 32);
        if (chunk < 1) chunk = 1;
        if (chunk > n_frames) chunk = n_frames;
        const size_t ftl = (size_t)chunk * n_tiles * 32;
        const size_t need[N_DWORK] = {ftl * N * 2, ftl * N * 4, ftl * DO_N * 4, ftl * N * 4, N < 160 ? ftl * N * 4 : 0};
        for (int k = 0; k < N_DWORK; k++) if (need[k] && ensure(&c->work[k], &c->work_cap[k], need[k])) goto out;
        b.xq = (int16_t *)c->work[W_XQ]; b.spec = (int32_t *)c->work[W_SPEC]; b.dout = (int32_t *)c->work[W_DOUT];
        b.yrow = (int32_t *)c->work[W_YROW]; b.work = (int32_t *)c->work[W_WORK];
        const int n_chunks = (n_frames + chunk - 1) / chunk, nb0 = num_bytes[ROW(0)], chn0 = decs[0]->cfg.channels;
        const size_t in_pitch = n_streams > 1 ? in_stream_stride : (size_t)nb0 * n_frames;
        const size_t out_pitch = n_streams > 1 ? pcm_stream_stride : (size_t)chn0 * n_frames * N * sb;
        if (rect) {
            if (ensure_events(c, n_chunks)) goto out;
            if (dfail(cudaEventRecord(c->in_ev[0], c->s), "cudaEventRecord") || dfail(cudaStreamWaitEvent(c->cp, c->in_ev[0], 0), "cudaStreamWaitEvent")) goto out; /* after the tables */
            if (!in_dev && nb0 > 0)
                for (int k = 0; k < n_chunks; k++) {
                    const int f0 = k * chunk, nf = n_frames - f0 < chunk ? n_frames - f0 : chunk;
                    if (dfail(cudaMemcpy2DAsync((uint8_t *)c->d_in + (size_t)f0 * nb0, in_pitch, (const uint8_t *)in + (size_t)f0 * nb0, in_pitch,
                                                (size_t)nf * nb0, (size_t)n_streams, cudaMemcpyHostToDevice, c->cp), "payload upload") ||
                        dfail(cudaEventRecord(c->in_ev[k], c->cp), "cudaEventRecord")) goto out;
                }
        }
        for (int f0 = 0, k = 0; f0 < n_frames; f0 += chunk, k++) {
            b.frame0 = f0;
            b.n_frames = n_frames - f0 < chunk ? n_frames - f0 : chunk;
            if (rect && !in_dev && nb0 > 0 && dfail(cudaStreamWaitEvent(c->s, c->in_ev[k], 0), "cudaStreamWaitEvent")) goto out;
            if (c->prof && f0 == 0) cudaEventRecord(c->pev[0], c->s);
            int e = lc3d_launch_entropy(&b, c->s);
            if (c->prof && f0 == 0) cudaEventRecord(c->pev[1], c->s);
            if (!e) e = lc3d_launch_idct(&b, c->s);
            if (c->prof && f0 == 0) cudaEventRecord(c->pev[2], c->s);
            if (!e) e = lc3d_launch_synth(&b, c->s);
            if (c->prof && f0 == 0) cudaEventRecord(c->pev[3], c->s);
            if (e) { dfail((cudaError_t)e, "decoder kernel launch"); goto out; }
            c->launches += 3;
            if (rect && !out_dev) {
                if (dfail(cudaEventRecord(c->k_ev[k], c->s), "cudaEventRecord") || dfail(cudaStreamWaitEvent(c->dl, c->k_ev[k], 0), "cudaStreamWaitEvent")) goto out;
                for (int ch = 0; ch < chn0; ch++) {
                    const size_t off = ((size_t)ch * n_frames + f0) * N * sb;
                    if (dfail(cudaMemcpy2DAsync((uint8_t *)pcm + off, out_pitch, (const uint8_t *)d_pcm + off, out_pitch, (size_t)b.n_frames * N * sb,
                                                (size_t)n_streams, cudaMemcpyDeviceToHost, c->dl), "pcm download")) goto out;
                }
            }
        }
        if (!out_dev && !rect)
            for (int i = 0; i < n_streams; i++) {
                const size_t o = (size_t)ROW(i) * pcm_stream_stride, len = (size_t)decs[i]->cfg.channels * n_frames * N * sb;
                if (dfail(cudaMemcpyAsync((uint8_t *)pcm + o, (const uint8_t *)d_pcm + o, len, cudaMemcpyDeviceToHost, c->s), "pcm download")) goto out;
            }
        if (dfail(cudaMemcpyAsync(c->h_bfi, c->d_bfi, bfi_bytes, cudaMemcpyDeviceToHost, c->s), "flag download")) goto out;
        if (c->prof) cudaEventRecord(c->pev[4], c->s);
        if (dfail(cudaStreamSynchronize(c->s), "decoder batch") || dfail(cudaStreamSynchronize(c->dl), "decoder batch")) goto out;
        if (c->prof) {
            float ms;
            for (int k = 0; k < 3; k++) if (cudaEventElapsedTime(&ms, c->pev[k], c->pev[k + 1]) == cudaSuccess) c->prof_ms[k] += ms;
            if (cudaEventElapsedTime(&ms, c->pev[0], c->pev[4]) == cudaSuccess) c->prof_ms[3] += ms;
        }
        int any = 0;
        /* device order is [frame][lane]; the caller gets [stream][frame] (a stream is bad in a frame if any channel is) */
        if (bfi_out) for (int i = 0; i < n_streams; i++) memset(bfi_out + (size_t)ROW(i) * n_frames, 0, (size_t)n_frames);
        for (int f = 0; f < n_frames; f++) {
            int l2 = 0;
            for (int i = 0; i < n_streams; i++)
                for (int ch = 0; ch < decs[i]->cfg.channels; ch++, l2++)
                    if (c->h_bfi[(size_t)f * n_lanes + l2]) { any = 1; if (bfi_out) bfi_out[(size_t)ROW(i) * n_frames + f] = 1; }
        }
        ret = any ? LC3_DECODE_ERROR : LC3_OK;
    }
out:
    pthread_mutex_unlock(&c->lock);
    return ret;
#undef ROW
}

/* Decoders of different sample rates, frame durations or devices in one call: grouped (in order of first appearance) and
 * decoded group by group. */
static LC3_Error dec_batch(LC3_Dec *const *decs, int n_streams, int n_frames, const void *in, size_t in_stream_stride,
                           const int *num_bytes, void *pcm, size_t pcm_stream_stride, int bitdepth, uint8_t *bfi_out)
{
    if (!decs || !in || !pcm || !num_bytes) return LC3_NULL_ERROR;
    if (bitdepth != 16 && bitdepth != 24) return LC3_ERROR;
    if (n_streams <= 0 || n_frames <= 0) return LC3_OK;
    for (int i = 0; i < n_streams; i++) if (!dec_ok(decs[i])) return LC3_NULL_ERROR;
    pthread_mutex_lock(&g_dlock);
    const unsigned mark = ++g_dcall;
    pthread_mutex_unlock(&g_dlock);
    int uniform = 1;
    for (int i = 0; i < n_streams; i++) {
        if ((unsigned)decs[i]->batch_mark == mark) { lc3b_set_error("lc3plus_dec_batch: a decoder appears twice in one call"); return LC3_ERROR; }
        decs[i]->batch_mark = (int)mark;
        if (decs[i]->cfg.fs != decs[0]->cfg.fs || decs[i]->cfg.dms != decs[0]->cfg.dms || decs[i]->device != decs[0]->device) uniform = 0;
    }
    if (uniform) return dec_bucket(decs, NULL, n_streams, n_streams, n_frames, in, in_stream_stride, num_bytes, pcm, pcm_stream_stride, bitdepth, bfi_out);
    if (is_dev_ptr(in) || is_dev_ptr(pcm)) {
        for (int i = 1; i < n_streams; i++)
            if (decs[i]->device != decs[0]->device) { lc3b_set_error("lc3plus_dec_batch: device buffers with decoders of several devices: make one call per device"); return LC3_ERROR; }
    }
    LC3_Dec **list = (LC3_Dec **)malloc((size_t)n_streams * sizeof(*list));
    int *rows = (int *)malloc((size_t)n_streams * sizeof(int));
    unsigned char *done = (unsigned char *)calloc((size_t)n_streams, 1);
    if (!list || !rows || !done) { free(list); free(rows); free(done); return LC3_ERROR; }
    LC3_Error ret = LC3_OK;
    for (int i0 = 0; i0 < n_streams; i0++) {
        if (done[i0]) continue;
        int n = 0;
        for (int i = i0; i < n_streams; i++)
            if (!done[i] && decs[i]->cfg.fs == decs[i0]->cfg.fs && decs[i]->cfg.dms == decs[i0]->cfg.dms && decs[i]->device == decs[i0]->device) {
                list[n] = decs[i]; rows[n] = i; n++; done[i] = 1;
            }
        const LC3_Error r = dec_bucket(list, rows, n, n_streams, n_frames, in, in_stream_stride, num_bytes, pcm, pcm_stream_stride, bitdepth, bfi_out);
        if (r == LC3_DECODE_ERROR) ret = r;
        else if (r != LC3_OK) { ret = r; break; }
    }
    free(list); free(rows); free(done);
    return ret;
}

LC3_Error lc3plus_dec_batch(LC3_Dec *const *decoders, int n_streams, int n_frames, const void *in_bytes, size_t in_stream_stride_bytes,
                            const int *num_bytes, void *pcm, size_t pcm_stream_stride_bytes, int bitdepth, uint8_t *bfi_flags)
{
    return dec_batch(decoders, n_streams, n_frames, in_bytes, in_stream_stride_bytes, num_bytes, pcm, pcm_stream_stride_bytes, bitdepth, bfi_flags);
}

static LC3_Error dec_one(LC3_Dec *dec, void *input_bytes, int num_bytes, void **out, int bitdepth, void *scratch, int bfi_ext)
{ /* fx/lc3.c:351-360, fx/dec_lc3.c:325-487 for one frame */
    if (!dec_ok(dec) || !input_bytes || !out || !scratch) return LC3_NULL_ERROR;
    for (int ch = 0; ch < dec->cfg.channels; ch++) if (!out[ch]) return LC3_NULL_ERROR;
    if (bitdepth != 16 && bitdepth != 24) return LC3_ERROR;
    const int N = dec->cfg.N, sb = bitdepth == 16 ? 2 : 4, chn = dec->cfg.channels;
    int nb = bfi_ext ? 0 : num_bytes;
    unsigned char tmp[2 * LC3K_MAX_N * 4];
    LC3_Dec *list[1] = {dec};
    LC3_Error r = dec_batch(list, 1, 1, input_bytes, (size_t)(nb > 0 ? nb : 1), &nb, tmp, (size_t)chn * N * sb, bitdepth, NULL);
    if (r != LC3_OK && r != LC3_DECODE_ERROR) return r;
    for (int ch = 0; ch < chn; ch++) memcpy(out[ch], tmp + (size_t)ch * N * sb, (size_t)N * sb);
    return r;
}
LC3_Error lc3_dec16(LC3_Dec *dec, void *input_bytes, int num_bytes, int16_t **output_samples, void *scratch, int bfi_ext)
{
    return dec_one(dec, input_bytes, num_bytes, (void **)output_samples, 16, scratch, bfi_ext);
}
LC3_Error lc3_dec24(LC3_Dec *dec, void *input_bytes, int num_bytes, int32_t **output_samples, void *scratch, int bfi_ext)
{
    return dec_one(dec, input_bytes, num_bytes, (void **)output_samples, 24, scratch, bfi_ext);
}

/* NEW: channel-frames per chunk of the decoder's pipeline (0 restores the default) */
void lc3plus_dec_set_chunk_frames(int channel_frames) { g_dchunk = channel_frames > 0 ? channel_frames : 5 << 15; }
long long lc3plus_dec_kernel_launches(void)
{
    long long n = 0;
    for (int d = 0; d < LC3D_MAX_DEV; d++) if (g_dctx[d]) n += g_dctx[d]->launches;
    return n;
}
/* profiling: ms[4] = the entropy, DCT and synthesis kernels of the FIRST chunk of every call, and the device span of the whole call
 * (first kernel to the end of the downloads), accumulated */
void lc3plus_dec_profile_enable(int on)
{
    for (int d = 0; d < LC3D_MAX_DEV; d++) if (g_dctx[d]) { g_dctx[d]->prof = on; memset(g_dctx[d]->prof_ms, 0, sizeof(g_dctx[d]->prof_ms)); }
}
void lc3plus_dec_profile_read(double *ms)
{
    for (int k = 0; k < 4; k++) ms[k] = 0;
    for (int d = 0; d < LC3D_MAX_DEV; d++) if (g_dctx[d]) for (int k = 0; k < 4; k++) ms[k] += g_dctx[d]->prof_ms[k];
}
