/* Host-side decoder configuration: the arithmetic of fx/setup_dec_lc3.c:83-336 and the argument checks of
 * fx/lc3.c (decoder half), with no device dependency (plain C). */
#ifndef LC3_DEC_CONFIG_H
#define LC3_DEC_CONFIG_H
#include "lc3_dec_layout.h"

typedef struct {
    int fs_in, fs, fs_idx, channels, dms, dms_idx;
    int N, yLen, W_off, W_len, la_zeroes, ola_len, nbands, bands_off, bw_bits, ltpf_x_len, ltpf_y_len;
} LC3DCfg;

/* return values are LC3_Error codes (fx/lc3.h:102-126) */
int lc3dcfg_init(LC3DCfg *c, int samplerate, int channels);
int lc3dcfg_set_frame_dms(LC3DCfg *c, int dms);
/* update_dec_bitrate (fx/setup_dec_lc3.c:271-336) for a channel payload of nbytes: fills w[DCFG_N] */
int lc3dcfg_lane_words(const LC3DCfg *c, int nbytes, int32_t *w);
int lc3dcfg_delay(const LC3DCfg *c);
void lc3dcfg_fill_batch(const LC3DCfg *c, LC3D_Batch *b);
#endif
