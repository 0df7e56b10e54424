/* Host-side encoder configuration: the arithmetic of fx/setup_enc_lc3.c:38-281 and the argument checks of
 * fx/lc3.c:94-244, with no device dependency (plain C).  lc3_api.c keeps one LC3Cfg inside the caller's
 * LC3_Enc block and mirrors the per-channel words into the device-side configuration pool. */
#ifndef LC3_CONFIG_H
#define LC3_CONFIG_H
#include "lc3_layout.h"

typedef struct {
    int fs_in, fs, fs_idx, channels, dms, dms_idx;
    int N, yLen, W_off, W_len, la_zeroes, mdct_mem_len, nbands, bands_off, nsub, ltpf_mem_in_len, r12k8_mem_in_len, bw_bits;
    int bitrate, br_set, bandwidth, bw_cutoff_bin, bw_index;
    int epmode, epmr, ccc;  /* error protection mode 0..4, mode request 0..3, combined channel coding (derived) */
    int32_t ch[2][CFG_N];   /* per-channel CFG_* words */
    int attack_reset[2];    /* set when set_bitrate switched attack handling off (state must be cleared) */
} LC3Cfg;

/* return values are LC3_Error codes (fx/lc3.h:102-126) */
int lc3cfg_init(LC3Cfg *c, int samplerate, int channels);
int lc3cfg_set_frame_dms(LC3Cfg *c, int dms);
int lc3cfg_set_bitrate(LC3Cfg *c, int bitrate);
int lc3cfg_set_bandwidth(LC3Cfg *c, int bandwidth);
int lc3cfg_set_ep_mode(LC3Cfg *c, int epmode);
int lc3cfg_set_ep_mode_request(LC3Cfg *c, int epmr);
int lc3cfg_out_bytes(const LC3Cfg *c, int ch); /* bytes channel ch occupies in the output frame before channel ch+1 */
int lc3cfg_num_bytes(const LC3Cfg *c);
int lc3cfg_real_bitrate(const LC3Cfg *c);
void lc3cfg_fill_batch(const LC3Cfg *c, LC3K_Batch *b);
#endif
