/* ORACLE (test infrastructure, not product code): internal declarations of the plain-C
 * restatement of the ETSI LC3plus fixed-point encoder (reference: fx/ under
 * /root/reference/LC3plus_ETSI_src_v17171_20200723/src/fixed_point/).  Each function cites the
 * reference file:line it follows. */
#ifndef LC3O_INTERNAL_H
#define LC3O_INTERNAL_H
#include <string.h>
#include "lc3o_basop.h"

#define ROM_DECL(type, name, n) extern const type name[n]
#include "rom_tables_decl.inc"
#undef ROM_DECL

#define LC3O_MAX_N 480
#define LC3O_MAX_BANDS 64

/* ---- lc3o_math.c  (fx/basop_util.c) */
i32 o_log2(i32 x);                       /* BASOP_Util_Log2      :25-86 */
i32 o_invlog2(i32 x);                    /* BASOP_Util_InvLog2   :89-138 */
i16 o_isqrt16(i16 m, i16 *e);            /* ISqrt16              :212-234 */
i16 o_inv16(i16 m, i16 *e);              /* Inv16                :236-273 */
i32 o_isqrt32(i32 x, i16 *e);            /* Isqrt                :788-832 */
i16 o_log2_16(i32 x, i16 x_e);           /* BASOP_Util_Log2_16   :835-861 */
i16 o_invlog2_16(i16 x, i16 *y_e);       /* BASOP_Util_InvLog2_16:863-873 */
i16 o_headroom16(const i16 *x, int n, int zero_val); /* getScaleFactor16 / _0 :286-357 */
i16 o_headroom32(const i32 *x, int n, int zero_val); /* getScaleFactor32 / _0 :368-444 */
i16 o_div3216(i32 x, i16 y, i16 *s);     /* BASOP_Util_Divide3216_Scale :446-482 */
i16 o_div1616(i16 x, i16 y, i16 *s);     /* BASOP_Util_Divide1616_Scale :484-532 */
i32 o_norm32norm(const i32 *x, i16 headroom, i16 n, i16 *res_e); /* Norm32Norm :534-569 */
void o_scale_sig(i16 *x, int n, i16 e);  /* Scale_sig :571-595 */
i32 o_add_me(i32 am, i16 ae, i32 bm, i16 be, i16 *re); /* BASOP_Util_Add_Mant32Exp :633-684 */
i16 o_cmp_me(i32 am, i16 ae, i32 bm, i16 be);           /* BASOP_Util_Cmp_Mant32Exp :686-768 */

/* ---- lc3o_fft.c (fx/fft.c) : in-place complex FFT on interleaved re/im (stride 2) */
void o_cfft(i32 *x, int len, i16 *scale);

/* ---- lc3o_mdct.c (fx/mdct_fx.c, fx/dct4_fx.c) */
void o_mdct(const i16 *x, i16 x_exp, int N, const i16 *w, int wLen, i16 *mem, int memLen, i32 *y, i16 *y_e);
void o_dct4(i32 *p, i16 *p_e, int L);   /* dct_IV, fx/dct4_fx.c:16-138 (shared with the decoder's inverse MDCT) */

/* ---- lc3o_pitch.c (fx/resamp12k8_fx.c, fx/olpa_fx.c, fx/ltpf_coder_fx.c) */
typedef struct {
    i16 r12k8_mem_in[60];
    i32 r12k8_mem_50[2];
    i16 r12k8_mem_out[24];
    i16 olpa_mem_s12k8[3];
    i16 olpa_mem_s6k4[64 + 114 + 16];
    i16 olpa_mem_s6k4_exp, olpa_mem_pitch;
    i16 ltpf_mem_in[232 + 128 + 1];
    i16 ltpf_mem_in_exp, ltpf_mem_normcorr, ltpf_mem_mem_normcorr, ltpf_mem_ltpf_on, ltpf_mem_pitch;
} o_pitch_state;
void o_resamp12k8(const i16 *x, int x_len, o_pitch_state *st, int mem_in_len, i16 *y, int *y_len, int fs_idx, int dms);
void o_olpa(o_pitch_state *st, i16 *pitch, const i16 *s12k8, int len, i16 *normcorr, i16 s12k8_exp, int dms);
void o_ltpf(o_pitch_state *st, i16 *bits, i16 ol_pitch, i16 ltpf_enable, int mem_in_len, i16 param[3],
            const i16 *xin, int len, i16 ol_normcorr, i16 xin_exp, int dms);

/* ---- lc3o_spec.c (fx/attack_detector_fx.c, per_band_energy_fx.c, detect_cutoff_warped_fx.c) */
typedef struct {
    i16 filter_mem[2];
    i16 detected, position, scaling;
    i32 acc_energy;
} o_attdec_state;
void o_attack_detector(o_attdec_state *st, int attack_handling, const i16 *in, int N, int fs, i16 in_scaling);
void o_band_energy(i32 *en, i16 *en_exp, const i32 *d, i16 d_exp, const i16 *bands, int fs_idx, int nbands, int dms_idx);
i16 o_bw_detect(const i32 *en, i16 en_exp, int fs_idx, int dms_idx);

/* ---- lc3o_sns.c (fx/sns_compute_scf_fx.c, sns_quantize_scf_fx.c, pvq_*_fx.c, dct2_fx.c,
 *                  sns_interpolate_scf_fx.c, mdct_shaping_fx.c) */
void o_sns_compute(i32 *en, i16 en_exp, int fs_idx, int nbands, i16 *scf, int attack, i16 att_damping);
void o_sns_quantize(const i16 *scf, i32 idx[7], i16 *scf_q);
void o_sns_interp(const i16 *scf_q, i16 *g, i16 *g_exp, int nbands);
void o_sns_interp_dir(const i16 *scf_q, i16 *g, i16 *g_exp, int nbands, int inv_scf); /* inv_scf 0: the decoder side */
void o_idct16(const i16 *in, i16 *out);                 /* fx/dct2_fx.c:111-206 */
void o_pvq_unit_norm(i16 *xq, const i16 *y, i16 kmax);  /* fx/pvq_dec_fx.c:67-117 */
void o_mdct_shape(i32 *d, const i16 *g, const i16 *g_exp, const i16 *bands, int nbands);

/* ---- lc3o_tns.c (fx/tns_coder_fx.c, fx/levinson_fx.c) */
void o_tns(i16 *bits, i16 idx[16], i32 *x, int bw, i16 order[2], i16 *nfilt, int lpc_weighting, int nsub, int dms_idx);

/* ---- lc3o_rate.c (fx/estimate_global_gain_fx.c, quantize_spec_fx.c, adjust_global_gain_fx.c,
 *                   residual_coding_fx.c, noise_factor_fx.c, ari_codec.c:1269-1308) */
void o_est_gain(const i32 *x, i16 x_e, int lg, i16 nbitsSQ, i16 *gain, i16 *gain_e, i16 *qgain, i16 *qgain_min,
                i16 qgain_off, i32 *tbits_off, i16 *old_tbits, i16 old_specbits);
void o_quantize(const i32 *x, i16 x_e, i16 gain, i16 gain_e, i16 *xq, int nt, i16 target, i16 totalBits,
                i16 *nBits, i16 *nBits2, int fs_idx, i16 *lastnz, i16 *cdata, i16 *lsbMode, int mode);
void o_adjust_gain(i16 *gg, i16 gg_min, i16 gg_off, i16 *gain, i16 *gain_e, i16 target, i16 nBits, i16 *change, int fs_idx);
void o_residual(i16 x_e, const i32 *x, const i16 *xq, i16 gain, i16 gain_e, int L, i16 target, i16 nBits, uint8_t *res, i16 *nres);
i16 o_noise_factor(i16 x_e, const i32 *x, const i16 *xq, i16 gg, i16 gg_e, int bw, int dms, i16 target_bytes);

/* ---- lc3o_bits.c (fx/enc_entropy.c, fx/ari_codec.c encoder half) */
void o_side_info(uint8_t *bytes, i16 *bp_side, i16 *mask_side, i16 nbbits, i16 targetBytes, int L, i16 bw_bits,
                 i16 nfilt, i16 lsbMode, i16 lastnz, const i16 *tns_order, i16 fac_ns, i16 gg, i16 bw,
                 const i16 *ltpf_idx, const i32 *scf_idx, int bfi_ext, int fs_idx);
int o_ari_encode(uint8_t *bytes, i16 bp_side, i16 mask_side, i16 nbbits, const i16 *xq, const i16 *tns_order,
                 i16 nfilt, const i16 *tns_idx, i16 lastnz, const i16 *cdata, const uint8_t *res, i16 nres,
                 i16 lsbMode, int lpc_weighting, int n_pc);

/* ---- lc3o_fec.c (fx/al_fec.c encoder half, fx/reorder_bitstream_fx.c) */
int o_fec_n_pccw(int slot_bytes, int mode, int ccc);
int o_fec_data_size(int mode, int ccc, int slot_bytes);
int o_fec_n_pc(int mode, int n_pccw, int slot_bytes);
void o_fec_encode(int mode, int epmr, uint8_t *iobuf, int data_bytes, int slot_bytes, int n_pccw);
void o_reorder(uint8_t *bytes, int n_pccw, int n_pc, int b_left);
#endif
