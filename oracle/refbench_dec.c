/* ORACLE / BASELINE TOOLING (not product code): times the UNMODIFIED ETSI reference DECODER on host cores.
 * Built by oracle/Makefile against the reference's own sources (fx/lc3.h API) into oracle/_ref/refbench_dec.
 *
 *   refbench_dec <payloads.bin> <fs> <dms> <nbytes> <n_streams> <n_frames> <n_threads> [out.pcm|-] [repeat]
 *
 * payloads.bin holds uint8 [n_streams][n_frames][nbytes].  Each thread decodes a contiguous share of the streams in memory
 * (no file I/O inside the timed region), calling lc3_dec16 once per frame exactly as the reference CLI does
 * (fx/codec_exe.c:468).  The block of n_frames is decoded `repeat` times in a row (state carried).  Prints one line:
 * frames seconds frames_per_s threads bad_frames.  With out.pcm the int16 PCM [n_streams][n_frames*N] of the last repeat is
 * written after timing (used to cross-check the GPU path). */
#define _GNU_SOURCE
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include "lc3.h"

typedef struct { int lo, hi, bad; } span_t;
static int g_fs, g_dms, g_nb, g_nf, g_N, g_rep = 1;
static uint8_t *g_in;
static int16_t *g_out;
static pthread_barrier_t g_bar;

static double now(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return t.tv_sec + 1e-9 * t.tv_nsec; }

static void *worker(void *arg)
{
    span_t *sp = (span_t *)arg;
    int n = sp->hi - sp->lo;
    LC3_Dec **dec = (LC3_Dec **)malloc(sizeof(*dec) * (n > 0 ? n : 1));
    void *scratch = NULL;
    for (int s = 0; s < n; s++) {
        dec[s] = (LC3_Dec *)malloc(lc3_dec_get_size(g_fs, 1, LC3_PLC_ADVANCED));
        lc3_dec_init(dec[s], g_fs, 1, LC3_PLC_ADVANCED);
        lc3_dec_set_frame_dms(dec[s], g_dms);
        if (!scratch) scratch = malloc(lc3_dec_get_scratch_size(dec[s]));
    }
    int16_t tmp[480];
    pthread_barrier_wait(&g_bar);
    for (int r = 0; r < g_rep; r++)
    for (int f = 0; f < g_nf; f++)
        for (int s = 0; s < n; s++) {
            int16_t *out[1] = {g_out ? g_out + ((size_t)(sp->lo + s) * g_nf + f) * g_N : tmp};
            if (lc3_dec16(dec[s], g_in + ((size_t)(sp->lo + s) * g_nf + f) * g_nb, g_nb, out, scratch, 0)) sp->bad++;
        }
    for (int s = 0; s < n; s++) free(dec[s]);
    free(dec); free(scratch);
    return NULL;
}

int main(int argc, char **argv)
{
    if (argc < 8) { fprintf(stderr, "usage: refbench_dec payloads.bin fs dms nbytes n_streams n_frames n_threads [out.pcm|-] [repeat]\n"); return 2; }
    g_fs = atoi(argv[2]); g_dms = atoi(argv[3]); g_nb = atoi(argv[4]);
    int S = atoi(argv[5]); g_nf = atoi(argv[6]);
    int T = atoi(argv[7]);
    const char *outp = argc > 8 && strcmp(argv[8], "-") ? argv[8] : NULL;
    if (argc > 9) g_rep = atoi(argv[9]);
    g_N = (g_fs == 44100 ? 48000 : g_fs) / 100 * g_dms / 100;
    size_t nin = (size_t)S * g_nf * g_nb;
    g_in = (uint8_t *)malloc(nin);
    FILE *f = fopen(argv[1], "rb");
    if (!f || fread(g_in, 1, nin, f) != nin) { fprintf(stderr, "cannot read %s\n", argv[1]); return 1; }
    fclose(f);
    if (outp) g_out = (int16_t *)calloc((size_t)S * g_nf * g_N, 2);
    if (T > S) T = S;
    pthread_t *th = (pthread_t *)malloc(sizeof(*th) * T);
    span_t *sp = (span_t *)calloc(T, sizeof(*sp));
    pthread_barrier_init(&g_bar, NULL, T + 1);
    for (int t = 0; t < T; t++) {
        sp[t].lo = (int)((long long)S * t / T); sp[t].hi = (int)((long long)S * (t + 1) / T);
        pthread_create(&th[t], NULL, worker, &sp[t]);
    }
    pthread_barrier_wait(&g_bar);
    double t0 = now();
    int bad = 0;
    for (int t = 0; t < T; t++) { pthread_join(th[t], NULL); bad += sp[t].bad; }
    double dt = now() - t0;
    double frames = (double)S * g_nf * g_rep;
    printf("%.0f %.6f %.1f %d %d\n", frames, dt, frames / dt, T, bad);
    if (outp) {
        f = fopen(outp, "wb");
        if (!f || fwrite(g_out, 2, (size_t)S * g_nf * g_N, f) != (size_t)S * g_nf * g_N) { fprintf(stderr, "cannot write %s\n", outp); return 1; }
        fclose(f);
    }
    return 0;
}
