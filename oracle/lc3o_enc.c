/* ORACLE (test infrastructure, not product code).
 * Encoder setup and per-frame driver: restates fx/setup_enc_lc3.c:38-281 (configuration derived from
 * sample rate / frame size / bitrate), fx/lc3.c:124-267 (argument checks and error order),
 * fx/scale_signal24_fx.c:13-92 (24-bit input block scaling) and fx/enc_lc3.c:15-315 (stage order).
 * Error protection (epmode 1-4): sizes and the channel coder are in lc3o_fec.c.
 *
 * The oracle's own API (lc3o_*) is what tests/ call; it is deliberately not the product's C ABI.
 */
#include <stdlib.h>
#include "lc3o_internal.h"
#include "lc3o.h"

typedef struct {
    i16 mdct_mem[300];
    i32 mdct_mem32[300];
    i32 resamp_mem32[60];
    i32 tbits_off;
    i16 target_bytes, total_bits, tbits_init, tbits_ari, lpc_weighting, ltpf_enable, qgain_off;
    i16 mem_tbits, mem_specbits, x_exp, resamp_exp, attack_handling;
    i16 n_pccw, n_pc; /* partial concealment: codewords / nibbles (error protection modes 3 and 4) */
    o_attdec_state att;
    o_pitch_state pitch;
} chan_t;

struct lc3o_enc {
    int fs, fs_in, fs_idx, bitrate, channels, dms, dms_idx, br_set;
    int N, yLen, W_size, la_zeroes, mdct_mem_len, nbands, nsub, ltpf_mem_in_len, bw_bits, r12k8_mem_in_len;
    int bandwidth, bw_cutoff_bin, bw_index;
    int epmode, epmr, ccc; /* error protection mode, mode request, combined channel coding */
    const i16 *W, *bands;
    chan_t ch[2];
};

static int fs_ok(int fs) { return fs == 8000 || fs == 16000 || fs == 24000 || fs == 32000 || fs == 44100 || fs == 48000; }

/* fx/setup_enc_lc3.c:69-127 */
static void set_frame_params(lc3o_enc *e)
{
    int k = e->dms_idx * 5 + e->fs_idx;
    e->N = (e->fs / 100) >> e->dms_idx;
    e->bw_bits = rom_bw_cutoff_bits[e->fs_idx];
    e->la_zeroes = rom_win_laz[k];
    e->mdct_mem_len = e->N - e->la_zeroes;
    e->nsub = e->dms == 100 ? 3 : 2;
    e->ltpf_mem_in_len = e->dms == 25 ? 232 + 32 : 232;
    e->yLen = (400 >> e->dms_idx) < e->N ? (400 >> e->dms_idx) : e->N;
    e->W = &rom_win[rom_win_off[k]];
    e->W_size = rom_win_len[k];
    e->nbands = rom_bands_number[k];
    e->bands = &rom_bands_offset[rom_bands_offset_off[k]];
}

/* fx/setup_enc_lc3.c:130-281 */
static int update_bitrate(lc3o_enc *e, int bitrate)
{
    int minBR = (int)(20 * 8.0 * (10000.0 / e->dms) * (e->fs_in == 44100 ? 441. / 480 : 1));
    int maxBR = (int)(400 * 8.0 * (10000.0 / e->dms) * (e->fs_in == 44100 ? 441. / 480 : 1));
    minBR *= e->channels;
    maxBR *= e->channels;
    if (bitrate < minBR || bitrate > maxBR) return LC3O_BITRATE_ERROR;
    e->bitrate = bitrate;
    e->br_set = 1;
    const int slot_total = (int)((int32_t)bitrate * e->N / (8 * e->fs_in));
    e->ccc = e->channels > 1 && e->epmode && bitrate * e->N / (8 * e->fs_in) <= 160;
    if (e->epmode > 0) {
        int mb = bitrate * e->N / (8 * e->fs_in * e->channels);
        if (mb < 40 || mb > 300) return LC3O_BITRATE_ERROR; /* FEC_SLOT_BYTES_MIN / _MAX */
    }
    int total = slot_total;
    if (e->ccc) {
        total = o_fec_data_size(e->epmode, 1, slot_total);
        e->ch[0].n_pccw = (i16)o_fec_n_pccw(slot_total, e->epmode, 1);
        e->ch[0].n_pc = (i16)o_fec_n_pc(e->epmode, e->ch[0].n_pccw, slot_total);
    }
    for (int c = 0; c < e->channels; c++) {
        chan_t *s = &e->ch[c];
        const int channel_bytes = total / e->channels + (c < (total % e->channels));
        if (e->ccc) {
            s->target_bytes = (i16)channel_bytes;
        } else {
            s->target_bytes = (i16)o_fec_data_size(e->epmode, 0, channel_bytes);
            s->n_pccw = (i16)o_fec_n_pccw(channel_bytes, e->epmode, 0);
            s->n_pc = (i16)o_fec_n_pc(e->epmode, s->n_pccw, channel_bytes);
        }
        if (s->target_bytes < 20 || s->target_bytes > 435) return LC3O_BITRATE_ERROR;
        s->total_bits = s_shl(s->target_bytes, 3);
        s->tbits_init = s_sub(s->total_bits, (i16)(38 + 8 + 3 + e->bw_bits));
        s->tbits_init = s_sub(s->tbits_init, s_sub(17, s_norm((i16)(e->yLen - 1))));
        if (s->total_bits > 1280) s->tbits_init = s_sub(s->tbits_init, 1);
        if (s->total_bits > 2560) s->tbits_init = s_sub(s->tbits_init, 1);
        s->tbits_ari = s->total_bits;
        i16 thr = s_add(560, s_imult(80, (i16)e->fs_idx));
        if (e->dms == 25) {
            s->ltpf_enable = s_sub(lo16(l_shr(l_mult0(9830, s->total_bits), 12)), thr) < 0;
            s->lpc_weighting = 0;
        } else if (e->dms == 50) {
            s->ltpf_enable = s_sub(s_sub(s_imult(s->total_bits, 2), 160), thr) < 0;
            s->lpc_weighting = s->total_bits < 240;
        } else {
            s->lpc_weighting = s->total_bits < 480;
            s->ltpf_enable = s_sub(s->total_bits, thr) < 0;
        }
        int q = s->total_bits / (10 * (e->fs_idx + 1));
        s->qgain_off = (i16)-((q < 115 ? q : 115) + 105 + 5 * (e->fs_idx + 1));
        if (e->dms == 100 && ((e->fs_in >= 44100 && s->target_bytes >= 100) || (e->fs_in == 32000 && s->target_bytes >= 81)) &&
            s->target_bytes < 340) {
            s->attack_handling = 1;
        } else { /* attack detector is reset whenever it is switched off */
            s->attack_handling = 0;
            memset(&s->att, 0, sizeof(s->att));
        }
    }
    return LC3O_OK;
}

int lc3o_enc_size(void) { return (int)sizeof(lc3o_enc); }

/* fx/lc3.c:124-131 + fx/setup_enc_lc3.c:38-66 */
int lc3o_enc_init(lc3o_enc *e, int samplerate, int channels)
{
    if (!e) return LC3O_NULL_ERROR;
    if (!fs_ok(samplerate)) return LC3O_SAMPLERATE_ERROR;
    if (channels < 1 || channels > 2) return LC3O_CHANNELS_ERROR;
    memset(e, 0, sizeof(*e));
    e->fs = samplerate == 44100 ? 48000 : samplerate;
    e->fs_in = samplerate;
    e->fs_idx = e->fs / 10000;
    e->channels = channels;
    e->dms = 100;
    e->dms_idx = 0;
    e->r12k8_mem_in_len = lo16(l_shr(mpy32x16(e->fs, 20972), 9));
    for (int c = 0; c < channels; c++) { e->ch[c].x_exp = 15; e->ch[c].resamp_exp = 17; }
    set_frame_params(e);
    return LC3O_OK;
}

lc3o_enc *lc3o_enc_new(int samplerate, int channels)
{
    lc3o_enc *e = (lc3o_enc *)malloc(sizeof(lc3o_enc));
    if (e && lc3o_enc_init(e, samplerate, channels) != LC3O_OK) { free(e); e = NULL; }
    return e;
}
void lc3o_enc_free(lc3o_enc *e) { free(e); }

/* fx/lc3.c:215-223 */
int lc3o_enc_set_frame_dms(lc3o_enc *e, int dms)
{
    if (!e) return LC3O_NULL_ERROR;
    if (dms != 25 && dms != 50 && dms != 100) return LC3O_FRAMEMS_ERROR;
    if (e->br_set) return LC3O_BITRATE_SET_ERROR;
    e->dms = dms;
    e->dms_idx = dms == 100 ? 0 : (dms == 50 ? 1 : 2);
    set_frame_params(e);
    return LC3O_OK;
}

/* fx/lc3.c:186-191 */
int lc3o_enc_set_bitrate(lc3o_enc *e, int bitrate)
{
    if (!e) return LC3O_NULL_ERROR;
    if (bitrate <= 0) return LC3O_BITRATE_ERROR;
    return update_bitrate(e, bitrate);
}

/* fx/lc3.c:199-214 */
int lc3o_enc_set_ep_mode(lc3o_enc *e, int epmode)
{
    if (!e) return LC3O_NULL_ERROR;
    if ((unsigned)epmode > 4) return 9; /* LC3_EPMODE_ERROR */
    e->epmode = epmode;
    return e->br_set ? update_bitrate(e, e->bitrate) : LC3O_OK;
}
int lc3o_enc_set_ep_mode_request(lc3o_enc *e, int epmr)
{
    if (!e) return LC3O_NULL_ERROR;
    if ((unsigned)epmr > 3) return 12; /* LC3_EPMR_ERROR */
    e->epmr = epmr;
    return LC3O_OK;
}

/* fx/lc3.c:226-244 */
int lc3o_enc_set_bandwidth(lc3o_enc *e, int bandwidth)
{
    if (!e) return LC3O_NULL_ERROR;
    if (e->bandwidth != bandwidth) {
        int eff = e->fs_in > 40000 ? 40000 : e->fs_in;
        if (bandwidth * 2 > eff) return LC3O_BW_WARNING;
        e->bandwidth = bandwidth;
        e->bw_cutoff_bin = l_div(l_mult0((i16)bandwidth, (i16)e->dms), 2500);
        e->bw_index = s_max(0, (i16)(l_div(bandwidth, 2000) - 1));
    }
    return LC3O_OK;
}

int lc3o_enc_get_input_samples(const lc3o_enc *e) { return e ? e->N : 0; }
int lc3o_enc_get_num_bytes(const lc3o_enc *e) { return e ? (int)((int32_t)e->bitrate * e->N / (8 * e->fs_in)) : 0; }
int lc3o_enc_get_delay(const lc3o_enc *e) { return e ? e->N - 2 * e->la_zeroes : 0; }
int lc3o_enc_get_real_bitrate(const lc3o_enc *e)
{ /* fx/lc3.c:163-184 */
    if (!e) return 0;
    if (!e->br_set) return LC3O_BITRATE_UNSET_ERROR;
    int tb = 0;
    for (int c = 0; c < e->channels; c++) tb += e->ch[c].target_bytes;
    int br = (int)((tb * 80000.0 + e->dms - 1) / e->dms);
    if (e->fs_in == 44100) { int rem = br % 480; br = ((br - rem) / 480) * 441 + (rem * 441) / 480; }
    return br;
}

/* fx/scale_signal24_fx.c:13-92 */
static void scale24(lc3o_enc *e, chan_t *s, const int32_t *x, i16 *xs)
{
    int N = e->N, ml = e->mdct_mem_len, rl = e->r12k8_mem_in_len;
    i16 sc[6];
    sc[0] = s_sub(23, o_headroom32(x, N, 32));
    sc[1] = s_sub(23, o_headroom32(s->mdct_mem32, ml, 32));
    sc[2] = s_sub(23, o_headroom32(s->resamp_mem32, rl, 32));
    sc[3] = s_sub(s_sub(s->resamp_exp, 2), o_headroom32(s->pitch.r12k8_mem_50, 2, 32));
    sc[4] = s_sub(s_sub(s->resamp_exp, 2), o_headroom16(s->pitch.r12k8_mem_out, 24, 16));
    sc[5] = s_sub(s_sub(s->resamp_exp, 2), o_headroom16(s->pitch.olpa_mem_s12k8, 3, 16));
    s->x_exp = 7;
    for (int i = 0; i < 6; i++) s->x_exp = s_max(s->x_exp, sc[i]);
    i16 sh = s_sub(23, s->x_exp);
    for (int i = 0; i < N; i++) xs[i] = l_round(l_shl(x[i], sh));
    for (int i = 0; i < ml; i++) s->mdct_mem[i] = l_round(l_shl(s->mdct_mem32[i], sh));
    for (int i = 0; i < rl; i++) s->pitch.r12k8_mem_in[i] = l_round(l_shl(s->resamp_mem32[i], sh));
    sh = s_sub(s_sub(s->resamp_exp, 2), s->x_exp);
    s->resamp_exp = s_add(s->x_exp, 2);
    if (sh) {
        for (int i = 0; i < 2; i++) s->pitch.r12k8_mem_50[i] = l_shl(s->pitch.r12k8_mem_50[i], sh);
        for (int i = 0; i < 24; i++) s->pitch.r12k8_mem_out[i] = s_shl(s->pitch.r12k8_mem_out[i], sh);
        for (int i = 0; i < 3; i++) s->pitch.olpa_mem_s12k8[i] = s_shl(s->pitch.olpa_mem_s12k8[i], sh);
    }
    memcpy(s->mdct_mem32, &x[N - ml], ml * sizeof(i32));
    memmove(s->resamp_mem32, &x[N - rl], rl * sizeof(i32));
}

/* fx/enc_lc3.c:15-268 : one channel, one frame */
static void encode_channel(lc3o_enc *e, chan_t *s, int bps, const void *in, uint8_t *bytes, int bfi_ext, lc3o_trace *tr)
{
    i32 d[LC3O_MAX_N], en[64], scf_idx[7];
    i16 xs[LC3O_MAX_N], s12k8[130], scf[16], scf_q[16], g[64], g_exp[64], xq[LC3O_MAX_N], cd[3 * LC3O_MAX_N / 2];
    i16 tns_idx[16], tns_order[2], ltpf_idx[3];
    uint8_t res[LC3O_MAX_N];
    i16 d_exp, en_exp, pitch, normcorr, ltpf_bits, nfilt, tns_bits, gain, gain_e, qgain, qgain_min;
    i16 nBits, lastnz, lsbMode, change, nres = 0, bp_side, mask_side, bw;
    int len12;

    memset(d, 0, sizeof(d)); memset(xq, 0, sizeof(xq)); memset(cd, 0, sizeof(cd));
    if (bps == 24) scale24(e, s, (const int32_t *)in, xs);
    else memcpy(xs, in, e->N * sizeof(i16));

    o_mdct(xs, s->x_exp, e->N, e->W, e->W_size, s->mdct_mem, e->mdct_mem_len, d, &d_exp);
    o_resamp12k8(xs, e->N, &s->pitch, e->r12k8_mem_in_len, s12k8, &len12, e->fs_idx, e->dms);
    o_olpa(&s->pitch, &pitch, s12k8, len12, &normcorr, s->resamp_exp, e->dms);
    o_ltpf(&s->pitch, &ltpf_bits, pitch, s->ltpf_enable, e->ltpf_mem_in_len, ltpf_idx, s12k8, len12, normcorr,
           s->resamp_exp, e->dms);
    o_attack_detector(&s->att, s->attack_handling, xs, e->N, e->fs, s_sub(s->x_exp, 15));
    o_band_energy(en, &en_exp, d, d_exp, e->bands, e->fs_idx, e->nbands, e->dms_idx);
    bw = e->fs_idx > 0 ? o_bw_detect(en, en_exp, e->fs_idx, e->dms_idx) : 0;
    if (tr) { memcpy(tr->mdct, d, sizeof(d)); tr->mdct_exp = d_exp; memcpy(tr->ener, en, sizeof(en)); tr->ener_exp = en_exp; }
    o_sns_compute(en, en_exp, e->fs_idx, e->nbands, scf, s->att.detected, 16384);
    o_sns_quantize(scf, scf_idx, scf_q);
    o_sns_interp(scf_q, g, g_exp, e->nbands);
    o_mdct_shape(d, g, g_exp, e->bands, e->nbands);
    if (e->bandwidth) { /* fx/cutoff_bandwidth.c:14-28 */
        int b = e->bw_cutoff_bin;
        if (e->yLen > b) {
            for (int i = -1; i < 3; i++) d[b + i] = l_shr(d[b + i], (i16)(i + 2));
            for (int i = b + 3; i < e->yLen; i++) d[i] = 0;
        }
        bw = s_min(bw, (i16)e->bw_index);
    }
    o_tns(&tns_bits, tns_idx, d, bw, tns_order, &nfilt, s->lpc_weighting, e->nsub, e->dms_idx);

    i16 tbits_quant = s_sub(s->tbits_init, s_add(tns_bits, ltpf_bits));
    o_est_gain(d, d_exp, e->yLen, tbits_quant, &gain, &gain_e, &qgain, &qgain_min, s->qgain_off, &s->tbits_off,
               &s->mem_tbits, s->mem_specbits);
    o_quantize(d, d_exp, gain, gain_e, xq, e->yLen, tbits_quant, s->tbits_ari, &s->mem_specbits, &nBits, e->fs_idx,
               &lastnz, cd, &lsbMode, -1);
    o_adjust_gain(&qgain, qgain_min, s->qgain_off, &gain, &gain_e, tbits_quant, s->mem_specbits, &change, e->fs_idx);
    if (change == 1)
        o_quantize(d, d_exp, gain, gain_e, xq, e->yLen, tbits_quant, s->tbits_ari, NULL, &nBits, e->fs_idx, &lastnz,
                   cd, &lsbMode, 0);
    if (lsbMode == 0) o_residual(d_exp, d, xq, gain, gain_e, e->yLen, tbits_quant, nBits, res, &nres);
    i16 fac_ns = o_noise_factor(d_exp, d, xq, gain, gain_e, bw, e->dms, s->target_bytes);

    o_side_info(bytes, &bp_side, &mask_side, s->tbits_ari, s->target_bytes, e->yLen, (i16)e->bw_bits, nfilt, lsbMode,
                lastnz, tns_order, fac_ns, qgain, bw, ltpf_idx, scf_idx, bfi_ext, e->fs_idx);
    const int n_pc = (!e->ccc && s->n_pc > 0) ? s->n_pc : 0; /* fx/enc_lc3.c:247-259 */
    const int b_left = o_ari_encode(bytes, bp_side, mask_side, s->tbits_ari, xq, tns_order, nfilt, tns_idx, lastnz, cd, res, nres,
                                    lsbMode, s->lpc_weighting, n_pc);
    if (n_pc > 0) o_reorder(bytes, s->n_pccw, s->n_pc, b_left);
    if (tr) {
        memcpy(tr->spec, d, sizeof(d)); memcpy(tr->xq, xq, sizeof(xq)); memcpy(tr->scf_idx, scf_idx, sizeof(scf_idx));
        tr->bw = bw; tr->tns_bits = tns_bits; tr->ltpf_bits = ltpf_bits; tr->qgain = qgain; tr->lastnz = lastnz;
        tr->lsbMode = lsbMode; tr->fac_ns = fac_ns; tr->nBits = nBits; tr->attack = s->att.detected; tr->pitch = pitch;
    }
}

/* fx/lc3.c:246-267 + fx/enc_lc3.c:270-315 */
static int encode(lc3o_enc *e, void **in, int bps, uint8_t *out, int *num_bytes, lc3o_trace *tr)
{
    if (!e || !in || !out || !num_bytes) return LC3O_NULL_ERROR;
    for (int c = 0; c < e->channels; c++) if (!in[c]) return LC3O_NULL_ERROR;
    if (!e->br_set) return LC3O_BITRATE_UNSET_ERROR;
    int bfi_ext = *num_bytes == -1, n = 0;
    const int total = (int)((int32_t)e->bitrate * e->N / (8 * e->fs_in));
    for (int c = 0; c < e->channels; c++) {
        encode_channel(e, &e->ch[c], bps, in[c], out + n, bfi_ext, tr);
        if (e->epmode && !e->ccc) { /* one channel-coded slot per channel */
            const int slot = total / e->channels + (c < (total % e->channels));
            o_fec_encode(e->epmode, e->epmr, out + n, e->ch[c].target_bytes, slot, e->ch[c].n_pccw);
            n += slot;
        } else {
            n += e->ch[c].target_bytes;
        }
    }
    if (e->epmode > 0 && e->ccc) { /* both channels in one slot */
        o_fec_encode(e->epmode, e->epmr, out, n, total, e->ch[0].n_pccw);
        n = total;
    }
    *num_bytes = n;
    return LC3O_OK;
}
int lc3o_enc16(lc3o_enc *e, int16_t **in, uint8_t *out, int *num_bytes) { return encode(e, (void **)in, 16, out, num_bytes, NULL); }
int lc3o_enc24(lc3o_enc *e, int32_t **in, uint8_t *out, int *num_bytes) { return encode(e, (void **)in, 24, out, num_bytes, NULL); }
int lc3o_enc16_trace(lc3o_enc *e, int16_t **in, uint8_t *out, int *num_bytes, lc3o_trace *tr) { return encode(e, (void **)in, 16, out, num_bytes, tr); }

/* Convenience for tests and the CPU-baseline leg of bench.py: encode n_frames consecutive frames of
 * one mono/stereo stream held as planar int16 (pcm[ch][n_frames*N]) into out (frame-major). */
int lc3o_encode_stream16(lc3o_enc *e, const int16_t *pcm, long ch_stride, int n_frames, uint8_t *out)
{
    int nb = lc3o_enc_get_num_bytes(e), N = e->N;
    for (int f = 0; f < n_frames; f++) {
        int16_t *in[2] = {(int16_t *)pcm + (long)f * N, (int16_t *)pcm + ch_stride + (long)f * N};
        int n = 0, r = lc3o_enc16(e, in, out + (long)f * nb, &n);
        if (r) return r;
    }
    return LC3O_OK;
}
