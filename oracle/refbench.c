/* ORACLE / BASELINE TOOLING (not product code): times the UNMODIFIED ETSI reference encoder on host cores.
 * Built by oracle/Makefile against the reference's own sources (fx/lc3.h API) into oracle/_ref/refbench.
 *
 *   refbench <pcm.raw> <fs> <dms> <bitrate> <n_streams> <n_frames>  <n_threads> [out.bin|-] [repeat]
 *
 * pcm.raw holds int16 [n_streams][n_frames*N].  Each thread encodes a contiguous share of the streams in
 * memory (no file I/O inside the timed region), calling lc3_enc16 once per frame exactly as the reference
 * CLI does (fx/codec_exe.c:418).  Prints one line: frames seconds frames_per_s threads.  With out.bin the
 * block of n_frames is encoded `repeat` times in a row (state carried) to lengthen the run without a larger
 * file; with out.bin the bitstreams [n_streams][n_frames][nbytes] of the last repeat are written after timing (used to cross-check the GPU path).
 *
 *   refbench @spec.txt <pcm.raw> <n_frames> <n_threads> [out.bin|-] [repeat]
 *
 * General form (stereo, 24-bit input, mixed configurations, bitrate switching between passes as fx/codec_exe.c:340-349 does
 * between frames).  spec.txt: first line "S in_stride_bytes out_stride_bytes", then one line per stream
 * "fs dms channels bits n_bitrates br0 br1 ...": pass r (0-based, `repeat` passes over the n_frames) is encoded at bitrate
 * br[r % n_bitrates], set with lc3_enc_set_bitrate before the pass.  pcm.raw: stream s starts at s*in_stride_bytes, channels
 * planar, n_frames*N samples each (int16, or int32 for bits 24).  out.bin: frame f of stream s of the LAST pass at
 * s*out_stride_bytes + f*lc3_enc_get_num_bytes.  Prints channel-frames seconds channel-frames/s threads. */
#define _GNU_SOURCE
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include "lc3.h"

typedef struct { int lo, hi; } span_t;
static int g_fs, g_dms, g_br, g_nf, g_N, g_nb, g_rep = 1;
static int16_t *g_pcm;
static uint8_t *g_out;
static pthread_barrier_t g_bar;

static double now(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec + 1e-9 * t.tv_nsec; }

static void *worker(void *arg)
{
    span_t *sp = (span_t *)arg;
    int n = sp->hi - sp->lo;
    LC3_Enc **enc = (LC3_Enc **)malloc(sizeof(*enc) * (n > 0 ? n : 1));
    void *scratch = NULL;
    for (int s = 0; s < n; s++) {
        enc[s] = (LC3_Enc *)malloc(lc3_enc_get_size(g_fs, 1));
        lc3_enc_init(enc[s], g_fs, 1);
        lc3_enc_set_frame_dms(enc[s], g_dms);
        lc3_enc_set_bitrate(enc[s], g_br);
        if (!scratch) scratch = malloc(lc3_enc_get_scratch_size(enc[s]));
    }
    uint8_t tmp[LC3_MAX_BYTES];
    pthread_barrier_wait(&g_bar);
    double t0 = now();
    for (int r = 0; r < g_rep; r++)
    for (int f = 0; f < g_nf; f++)
        for (int s = 0; s < n; s++) {
            int16_t *in[1] = {g_pcm + ((size_t)(sp->lo + s) * g_nf + f) * g_N};
            int nb = 0;
            lc3_enc16(enc[s], in, g_out ? g_out + ((size_t)(sp->lo + s) * g_nf + f) * g_nb : tmp, &nb, scratch);
        }
    double t1 = now();
    for (int s = 0; s < n; s++) free(enc[s]);
    free(enc); free(scratch);
    double *res = (double *)malloc(2 * sizeof(double));
    res[0] = t0; res[1] = t1;
    return res;
}

/* ---- general form */
typedef struct { int fs, dms, ch, bits, nbr, br[16]; } spec_t;
static spec_t *g_spec;
static size_t g_istride, g_ostride;
static void *g_pcmv;

static void *worker_spec(void *arg)
{
    span_t *sp = (span_t *)arg;
    int n = sp->hi - sp->lo;
    LC3_Enc **enc = (LC3_Enc **)malloc(sizeof(*enc) * (n > 0 ? n : 1));
    void *scratch = malloc(LC3_ENC_MAX_SCRATCH_SIZE);
    for (int s = 0; s < n; s++) {
        const spec_t *c = &g_spec[sp->lo + s];
        enc[s] = (LC3_Enc *)malloc(lc3_enc_get_size(c->fs, c->ch));
        if (lc3_enc_init(enc[s], c->fs, c->ch) || lc3_enc_set_frame_dms(enc[s], c->dms)) { fprintf(stderr, "bad stream config\n"); exit(2); }
    }
    uint8_t tmp[LC3_MAX_BYTES];
    pthread_barrier_wait(&g_bar);
    double t0 = now();
    for (int r = 0; r < g_rep; r++) {
        for (int s = 0; s < n; s++) {
            const spec_t *c = &g_spec[sp->lo + s];
            if (r == 0 || c->nbr > 1)
                if (lc3_enc_set_bitrate(enc[s], c->br[r % c->nbr])) { fprintf(stderr, "bad bitrate %d\n", c->br[r % c->nbr]); exit(2); }
        }
        for (int f = 0; f < g_nf; f++)
            for (int s = 0; s < n; s++) {
                const spec_t *c = &g_spec[sp->lo + s];
                const int N = lc3_enc_get_input_samples(enc[s]), nb0 = lc3_enc_get_num_bytes(enc[s]), sb = c->bits == 16 ? 2 : 4;
                char *row = (char *)g_pcmv + (size_t)(sp->lo + s) * g_istride;
                void *in[2] = {row + (size_t)f * N * sb, row + ((size_t)g_nf + f) * N * sb};
                int nb = 0;
                uint8_t *o = g_out ? g_out + (size_t)(sp->lo + s) * g_ostride + (size_t)f * nb0 : tmp;
                if (c->bits == 16) lc3_enc16(enc[s], (int16_t **)in, o, &nb, scratch);
                else lc3_enc24(enc[s], (int32_t **)in, o, &nb, scratch);
            }
    }
    double t1 = now();
    for (int s = 0; s < n; s++) free(enc[s]);
    free(enc); free(scratch);
    double *res = (double *)malloc(2 * sizeof(double));
    res[0] = t0; res[1] = t1;
    return res;
}

static int main_spec(int argc, char **argv)
{
    if (argc < 5) { fprintf(stderr, "usage: refbench @spec.txt pcm.raw n_frames n_threads [out.bin|-] [repeat]\n"); return 2; }
    FILE *f = fopen(argv[1] + 1, "r");
    int S = 0;
    if (!f || fscanf(f, "%d %zu %zu", &S, &g_istride, &g_ostride) != 3 || S <= 0) { fprintf(stderr, "bad spec\n"); return 2; }
    g_spec = (spec_t *)calloc((size_t)S, sizeof(spec_t));
    double cf = 0;
    for (int s = 0; s < S; s++) {
        spec_t *c = &g_spec[s];
        if (fscanf(f, "%d %d %d %d %d", &c->fs, &c->dms, &c->ch, &c->bits, &c->nbr) != 5 || c->nbr < 1 || c->nbr > 16) { fprintf(stderr, "bad spec line %d\n", s); return 2; }
        for (int k = 0; k < c->nbr; k++) if (fscanf(f, "%d", &c->br[k]) != 1) return 2;
        cf += c->ch;
    }
    fclose(f);
    g_nf = atoi(argv[3]);
    int T = atoi(argv[4]);
    if (T < 1) T = 1;
    if (T > 256) T = 256;
    if (T > S) T = S;
    size_t n = (size_t)S * g_istride;
    g_pcmv = malloc(n);
    f = fopen(argv[2], "rb");
    if (!f || fread(g_pcmv, 1, n, f) != n) { fprintf(stderr, "cannot read %zu bytes from %s\n", n, argv[2]); return 2; }
    fclose(f);
    if (argc > 5 && strcmp(argv[5], "-")) g_out = (uint8_t *)calloc((size_t)S, g_ostride);
    if (argc > 6) g_rep = atoi(argv[6]) > 0 ? atoi(argv[6]) : 1;
    pthread_barrier_init(&g_bar, NULL, T);
    pthread_t th[256];
    span_t sp[256];
    for (int t = 0; t < T; t++) {
        sp[t].lo = (int)((long long)S * t / T); sp[t].hi = (int)((long long)S * (t + 1) / T);
        pthread_create(&th[t], NULL, worker_spec, &sp[t]);
    }
    double t0 = 1e300, t1 = 0;
    for (int t = 0; t < T; t++) {
        double *r; pthread_join(th[t], (void **)&r);
        if (r[0] < t0) t0 = r[0];
        if (r[1] > t1) t1 = r[1];
        free(r);
    }
    double frames = cf * g_nf * g_rep;
    printf("%.0f %.6f %.1f %d\n", frames, t1 - t0, frames / (t1 - t0), T);
    if (g_out) { f = fopen(argv[5], "wb"); fwrite(g_out, 1, (size_t)S * g_ostride, f); fclose(f); }
    return 0;
}

int main(int argc, char **argv)
{
    if (argc > 1 && argv[1][0] == '@') return main_spec(argc, argv);
    if (argc < 8) { fprintf(stderr, "usage: refbench pcm.raw fs dms bitrate n_streams n_frames n_threads [out.bin]\n"); return 2; }
    g_fs = atoi(argv[2]); g_dms = atoi(argv[3]); g_br = atoi(argv[4]);
    int S = atoi(argv[5]); g_nf = atoi(argv[6]);
    int T = atoi(argv[7]);
    if (T < 1) T = 1;
    if (T > 256) T = 256;
    if (T > S) T = S;
    {
        LC3_Enc *e = (LC3_Enc *)malloc(lc3_enc_get_size(g_fs, 1));
        if (lc3_enc_init(e, g_fs, 1) || lc3_enc_set_frame_dms(e, g_dms) || lc3_enc_set_bitrate(e, g_br)) { fprintf(stderr, "bad config\n"); return 2; }
        g_N = lc3_enc_get_input_samples(e); g_nb = lc3_enc_get_num_bytes(e);
        free(e);
    }
    size_t n = (size_t)S * g_nf * g_N;
    g_pcm = (int16_t *)malloc(n * 2);
    FILE *f = fopen(argv[1], "rb");
    if (!f || fread(g_pcm, 2, n, f) != n) { fprintf(stderr, "cannot read %zu samples from %s\n", n, argv[1]); return 2; }
    fclose(f);
    if (argc > 8 && strcmp(argv[8], "-")) g_out = (uint8_t *)malloc((size_t)S * g_nf * g_nb);
    if (argc > 9) g_rep = atoi(argv[9]) > 0 ? atoi(argv[9]) : 1;
    pthread_barrier_init(&g_bar, NULL, T);
    pthread_t th[256];
    span_t sp[256];
    for (int t = 0; t < T; t++) {
        sp[t].lo = (int)((long long)S * t / T); sp[t].hi = (int)((long long)S * (t + 1) / T);
        pthread_create(&th[t], NULL, worker, &sp[t]);
    }
    double t0 = 1e300, t1 = 0;
    for (int t = 0; t < T; t++) {
        double *r; pthread_join(th[t], (void **)&r);
        if (r[0] < t0) t0 = r[0];
        if (r[1] > t1) t1 = r[1];
        free(r);
    }
    double frames = (double)S * g_nf * g_rep;
    printf("%.0f %.6f %.1f %d\n", frames, t1 - t0, frames / (t1 - t0), T);
    if (g_out) { f = fopen(argv[8], "wb"); fwrite(g_out, 1, (size_t)S * g_nf * g_nb, f); fclose(f); }
    return 0;
}
