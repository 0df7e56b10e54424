/* See lc3_dec_config.h.  Reference: fx/setup_dec_lc3.c:83-336, fx/lc3.c:270-400. */
#include <string.h>
#include "lc3_dec_config.h"

#define ROM_TAB(type, name, n) static const type name[n] __attribute__((unused))
#include "rom_tables.inc"
#undef ROM_TAB

enum { E_OK = 0, E_NULL = 3, E_SAMPLERATE = 4, E_CHANNELS = 5, E_FRAMEMS = 10, E_NUMBYTES = 7 };

static int mpy_32_16(int32_t x, int y) { return (int)(((int64_t)x * y) >> 15); }

static void frame_params(LC3DCfg *c)
{ /* fx/setup_dec_lc3.c:169-231 */
    const int k = c->dms_idx * 5 + c->fs_idx;
    c->N = (c->fs / 100) >> c->dms_idx;
    c->la_zeroes = rom_win_laz[k];
    c->ola_len = c->N - c->la_zeroes;
    c->yLen = (400 >> c->dms_idx) < c->N ? (400 >> c->dms_idx) : c->N;
    c->W_off = rom_win_off[k];
    c->W_len = rom_win_len[k];
    c->nbands = rom_bands_number[k];
    c->bands_off = rom_bands_offset_off[k];
}

int lc3dcfg_init(LC3DCfg *c, int samplerate, int channels)
{ /* fx/lc3.c (lc3_dec_init), fx/setup_dec_lc3.c:83-106 */
    if (!c) return E_NULL;
    if (!(samplerate == 8000 || samplerate == 16000 || samplerate == 24000 || samplerate == 32000 ||
          samplerate == 44100 || samplerate == 48000)) return E_SAMPLERATE;
    if (channels < 1 || channels > 2) return E_CHANNELS;
    memset(c, 0, sizeof(*c));
    c->fs_in = samplerate;
    c->fs = samplerate == 44100 ? 48000 : samplerate;
    c->fs_idx = c->fs / 10000;
    c->channels = channels;
    c->dms = 100;
    c->dms_idx = 0;
    c->bw_bits = rom_bw_cutoff_bits[c->fs_idx];
    const int fmax = c->fs > 16000 ? c->fs : 16000;
    c->ltpf_x_len = (mpy_32_16(fmax, 16778) >> 11) - 2;
    c->ltpf_y_len = ((mpy_32_16(c->fs, 18678) - 1) >> 5) + 1 + (mpy_32_16(fmax, 16778) >> 12);
    frame_params(c);
    return E_OK;
}

int lc3dcfg_set_frame_dms(LC3DCfg *c, int dms)
{
    if (!c) return E_NULL;
    if (dms != 25 && dms != 50 && dms != 100) return E_FRAMEMS;
    c->dms = dms;
    c->dms_idx = dms == 100 ? 0 : (dms == 50 ? 1 : 2);
    frame_params(c);
    return E_OK;
}

int lc3dcfg_lane_words(const LC3DCfg *c, int nbytes, int32_t *w)
{ /* fx/setup_dec_lc3.c:271-336 */
    const int total_bits = nbytes << 3;
    int lpcw = total_bits < 480, tb = total_bits;
    int q = total_bits / (10 * (c->fs_idx + 1));
    if (q > 115) q = 115;
    const int qoff = -(q + 105 + 5 * (c->fs_idx + 1));
    const int tmp = 80 * c->fs_idx;
    if (c->dms == 25) { lpcw = 0; tb = (int)((19661 * (int32_t)total_bits) >> 13); tb = (int16_t)tb; }
    else if (c->dms == 50) { lpcw = total_bits < 240; tb = total_bits * 2 - 160; }
    if (nbytes < 20 || nbytes > 435) return E_NUMBYTES; /* MIN_NBYTES, MAX_NBYTES (fx/defines.h:141-142) */
    int sf;
    if (tb < 320 + tmp) sf = 0;
    else if (tb < 400 + tmp) sf = 1;
    else if (tb < 480 + tmp) sf = 2;
    else if (tb < 560 + tmp) sf = 3;
    else sf = -1;
    w[DCFG_BYTES] = nbytes; w[DCFG_LPC_WEIGHTING] = lpcw; w[DCFG_QGAIN_OFF] = qoff; w[DCFG_LTPF_SCALE_IDX] = sf;
    return E_OK;
}

int lc3dcfg_delay(const LC3DCfg *c)
{ /* lc3_dec_get_delay: frame_length - 2 * la_zeroes, in samples at the caller's rate */
    return c->N - 2 * c->la_zeroes;
}

void lc3dcfg_fill_batch(const LC3DCfg *c, LC3D_Batch *b)
{
    b->fs_idx = c->fs_idx; b->dms = c->dms; b->dms_idx = c->dms_idx; b->N = c->N; b->yLen = c->yLen; b->nbands = c->nbands;
    b->W_off = c->W_off; b->W_len = c->W_len; b->la_zeroes = c->la_zeroes; b->ola_len = c->ola_len; b->bands_off = c->bands_off;
    b->bw_bits = c->bw_bits; b->ltpf_x_len = c->ltpf_x_len; b->ltpf_y_len = c->ltpf_y_len;
}
