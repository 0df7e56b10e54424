/* ORACLE (test infrastructure, not product code): plain-C restatement of the ETSI LC3plus fixed-point DECODER for good
 * frames (fx/dec_lc3.c:15-322 with bfi = 0 and n_pc = 0: no concealment, no channel decoder), one frame of one channel at a
 * time, in the reference's own order and loop structure.  Each function cites the reference file:line it follows.
 *
 * Parity status: PINNED.  tests/test_oracle_dec.py checks it sample for sample against the unmodified reference decoder
 * compiled from /root/reference (oracle/_ref/liblc3ref.so) and against the committed digests of that decoder's output
 * (tests/golden/dec_golden.json).  A frame the reference would conceal returns LC3O_DECODE_ERROR and is synthesised from
 * a zero spectrum (the product's documented behaviour, include/lc3plus_b200.h). */
#include <stdlib.h>
#include "lc3o.h"
#include "lc3o_internal.h"

#define ROM_TAB(type, name, n) static const type name[n] __attribute__((unused))
#include "rom_dec_tables.inc"
#undef ROM_TAB

#define MAXLAG 8
typedef uint8_t u8;

struct lc3o_dec {
    int fs_in, fs, fs_idx, channels, dms, dms_idx;
    int N, yLen, W_off, W_len, la_zeroes, ola_len, nbands, bands_off, bw_bits, ltpf_x_len, ltpf_y_len;
    struct {
        i16 ola[LC3O_MAX_N], ola_e;
        i16 ltpf_x[10 + LC3O_MAX_N], ltpf_y[861 + LC3O_MAX_N], ltpf_e;
        i16 pitch_int, pitch_fr, gain, active, scale_idx;
    } ch[2];
};

/* ---- setup (fx/setup_dec_lc3.c:83-106, :169-231) */
static void frame_params(lc3o_dec *d)
{
    const int k = d->dms_idx * 5 + d->fs_idx;
    d->N = (d->fs / 100) >> d->dms_idx;
    d->la_zeroes = rom_win_laz[k];
    d->ola_len = d->N - d->la_zeroes;
    d->yLen = (400 >> d->dms_idx) < d->N ? (400 >> d->dms_idx) : d->N;
    d->W_off = rom_win_off[k];
    d->W_len = rom_win_len[k];
    d->nbands = rom_bands_number[k];
    d->bands_off = rom_bands_offset_off[k];
}
lc3o_dec *lc3o_dec_new(int samplerate, int channels)
{
    if (!(samplerate == 8000 || samplerate == 16000 || samplerate == 24000 || samplerate == 32000 || samplerate == 44100 || samplerate == 48000)) return NULL;
    if (channels < 1 || channels > 2) return NULL;
    lc3o_dec *d = (lc3o_dec *)calloc(1, sizeof(*d));
    if (!d) return NULL;
    d->fs_in = samplerate;
    d->fs = samplerate == 44100 ? 48000 : samplerate;
    d->fs_idx = d->fs / 10000;
    d->channels = channels;
    d->dms = 100;
    d->bw_bits = rom_bw_cutoff_bits[d->fs_idx];
    const i32 fmax = d->fs > 16000 ? d->fs : 16000;
    d->ltpf_x_len = lo16(l_shr(mpy32x16(fmax, 16778), 11)) - 2;
    d->ltpf_y_len = lo16(l_shr(mpy32x16(d->fs, 18678) - 1, 5)) + 1 + lo16(l_shr(mpy32x16(fmax, 16778), 12));
    for (int c = 0; c < 2; c++) d->ch[c].scale_idx = -1;
    frame_params(d);
    return d;
}
void lc3o_dec_free(lc3o_dec *d) { free(d); }
int lc3o_dec_set_frame_dms(lc3o_dec *d, int dms)
{
    if (!d) return LC3O_NULL_ERROR;
    if (dms != 25 && dms != 50 && dms != 100) return LC3O_FRAMEMS_ERROR;
    d->dms = dms;
    d->dms_idx = dms == 100 ? 0 : (dms == 50 ? 1 : 2);
    frame_params(d);
    return LC3O_OK;
}
int lc3o_dec_get_output_samples(const lc3o_dec *d) { return d ? d->N : 0; }
int lc3o_dec_get_delay(const lc3o_dec *d) { return d ? d->N - 2 * d->la_zeroes : 0; }

/* ---- side information (fx/ari_codec.c:1530-1556 read_bit, fx/dec_entropy.c:284-301 read_indice) */
typedef struct { const u8 *p; int n; i16 bp, mask; int oob; } side_rd;
static i16 rd_bit(side_rd *r)
{
    i16 bit = 0;
    if (r->bp < 0 || r->bp >= r->n) r->oob = 1;
    else if (r->p[r->bp] & r->mask) bit = 1;
    r->mask = (i16)(r->mask << 1);
    if (r->mask == 0x100) r->mask = 1;
    if (r->mask == 1) r->bp--;
    return bit;
}
static i16 rd_idx(side_rd *r, int numbits)
{
    i16 v = rd_bit(r);
    for (int i = 1; i < numbits; i++) v = s_add(v, s_lshl(rd_bit(r), i));
    return v;
}

typedef struct {
    i16 nfilt, lsbMode, lastnz, tns_order[2], fac_ns_idx, gg_idx, bw_idx, ltpf[3];
    i32 scf[7];
} side_t;

/* ac_dec_split_st2VQ_CW, fx/dec_entropy.c:303-369 */
static i16 split_cw(i32 cw, i32 szA, i32 szB, i32 *cwA, i32 *cwB, i16 *lsb)
{
    const i32 max_size = (i32)((u32)szB * (u32)szA);
    if (l_sub(cw, max_size) >= 0) { *cwA = 0; *cwB = 0; *lsb = 0; return 1; }
    i16 start = 0;
    if (l_sub(cw, l_shr(max_size, 1)) >= 0) start = (i16)l_shr(szB, 1);
    i16 ind = start;
    i32 t = l_neg(cw);
    t = l_add(t, (i32)((u32)start * (u32)szA));
    const i16 fin = s_add(start, (i16)l_shr(szB, 1));
    for (int i = start; i < fin; i++) {
        ind = s_add(ind, 1);
        t = l_add(t, szA);
        if (t > 0) ind = s_sub(ind, 1);
    }
    *cwB = ind;
    *cwA = l_sub(cw, (i32)((u32)ind * (u32)szA));
    *lsb = 0;
    *cwB = l_sub(*cwB, 2);
    if (*cwB < 0) *lsb = 1;
    *cwB = l_mac0(*cwB, 2, *lsb);
    return 0;
}

/* processDecoderEntropy_fx, fx/dec_entropy.c:22-194; returns bfi */
static int side_info(const lc3o_dec *d, side_rd *r, side_t *s)
{
    if (d->bw_bits > 0) {
        s->bw_idx = rd_idx(r, d->bw_bits);
        if (d->fs_idx < s->bw_idx) { s->bw_idx = (i16)d->fs_idx; return 1; }
    } else s->bw_idx = (i16)d->fs_idx;
    s->nfilt = (s->bw_idx >= 3 && d->dms >= 50) ? 2 : 1;
    const int L = 14 - s_norm(s_neg((i16)d->yLen));
    i16 n = rd_idx(r, L);
    n = s_add(n, 1);
    s->lastnz = s_shl(n, 1);
    if (s->lastnz > d->yLen) return 1;
    s->lsbMode = rd_bit(r);
    s->gg_idx = rd_idx(r, 8);
    s->tns_order[1] = 0;
    for (int k = 0; k < s->nfilt; k++) s->tns_order[k] = rd_bit(r);
    s->ltpf[0] = rd_idx(r, 1);
    s->scf[0] = rd_idx(r, 5);
    s->scf[1] = rd_idx(r, 5);
    const i16 msb = rd_bit(r);
    s->scf[2] = s_shl(msb, 1);
    s->scf[3] = rd_idx(r, rom_sns_gain_msb_bits[s->scf[2]]);
    s->scf[4] = rd_bit(r);
    i16 lsb, ber;
    if (msb == 0) {
        i32 cw = rd_idx(r, 13);
        cw |= (i32)rd_idx(r, 12) << 13;
        ber = split_cw(cw, (i32)rom_sns_mpvq_sz[0], (i32)(rom_sns_mpvq_sz[1] + rom_sns_mpvq_sz[3]), &s->scf[5], &s->scf[6], &lsb);
        if (lsb != 0) { s->scf[3] = l_add(l_shl(s->scf[3], 1), s->scf[6]); s->scf[6] = -2; }
    } else {
        i32 cw = rd_idx(r, 12);
        cw |= (i32)rd_idx(r, 12) << 12;
        s->scf[5] = cw;
        s->scf[6] = -1;
        lsb = 0; ber = 0;
        const i32 lim = l_add((i32)rom_sns_mpvq_sz[4], l_shl((i32)rom_sns_mpvq_sz[6], 1));
        if (l_sub(cw, lim) >= 0) ber = 1;
        else {
            cw = l_sub(cw, (i32)rom_sns_mpvq_sz[4]);
            if (cw >= 0) {
                lsb = 1;
                s->scf[3] = l_add(l_shl(s->scf[3], 1), cw & 1);
                s->scf[5] = l_shr(cw, 1);
                s->scf[6] = -2;
            }
        }
    }
    s->scf[2] = l_add(s->scf[2], lsb);
    if (ber > 0) return 1;
    if (s->ltpf[0] != 0) { s->ltpf[1] = rd_idx(r, 1); s->ltpf[2] = rd_idx(r, 9); }
    else { s->ltpf[1] = 0; s->ltpf[2] = 0; }
    s->fac_ns_idx = rd_idx(r, 3);
    return 0;
}

/* ---- arithmetic decoder (fx/ari_codec.c:1559-1803) */
typedef struct { u32 low, range, help; int ber, bp, n; const u8 *p; } ac_dec;
static u32 ac_byte(ac_dec *s)
{
    u32 v = 0;
    if (s->bp < s->n) v = s->p[s->bp]; else s->ber = 1;
    s->bp++;
    return v;
}
static void ac_init(ac_dec *s)
{
    s->low = 0; s->range = 0x00ffffff; s->ber = 0;
    for (int i = 0; i < 3; i++) s->low = (s->low << 8) + ac_byte(s);
}
static i16 ac_decode17(ac_dec *s, const uint16_t *cum)   /* ac_decode_fx :1589-1648 */
{
    s->help = s->range >> 10;
    i16 val = 0;
    if (s->low >= s->help * cum[8]) val = 8;
    if (s->low >= s->help * cum[val + 4]) val += 4;
    if (s->low >= s->help * cum[val + 2]) val += 2;
    if (s->low >= s->help * cum[val + 1]) {
        val += 1;
        if (val == 15) {
            if (s->low >= s->help * cum[16]) val = 16;
            if (s->low >= (s->help << 10)) s->ber = 1;
        }
    }
    return val;
}
static i16 ac_decode17s(ac_dec *s, const i16 *cum)       /* ac_decode_tns_coef :1697-1756 (signed table type) */
{
    uint16_t c[17];
    for (int i = 0; i < 17; i++) c[i] = (uint16_t)cum[i];
    return ac_decode17(s, c);
}
static i16 ac_decode_tns_order(ac_dec *s, const i16 *cum) /* :1651-1694 */
{
    s->help = s->range >> 10;
    i16 val = 0;
    if (s->low >= s->help * (u32)cum[4]) val = 4;
    if (s->low >= s->help * (u32)cum[val + 2]) val += 2;
    if (s->low >= s->help * (u32)cum[val + 1]) val += 1;
    if (s->low >= (s->help << 10)) s->ber = 1;
    return val;
}
static void ac_update(ac_dec *s, u32 cum, u32 freq)       /* ac_dec_update_fx :1758-1803 */
{
    s->low -= cum * s->help;
    s->range = s->help * freq;
    while (s->range < (1u << 16)) {
        s->low = ((s->low & 0xffff) << 8) + ac_byte(s);
        s->range <<= 8;
        if (s->ber) { s->range = 0x00ffffff; return; }
    }
}

typedef struct { i16 resBits, nf_seed, zero_frame, tns_idx[2 * MAXLAG]; u8 res[LC3O_MAX_N]; } ari_out;

/* processAriDecoder_fx, fx/ari_codec.c:409-1214 with n_pc = 0, mode 0; returns bfi */
static int ari_decode(const lc3o_dec *d, const u8 *bytes, int nbytes, side_rd *r, side_t *si, int lpc_weighting, i16 *x, ari_out *o)
{
    const i16 nbbits = (i16)(nbytes << 3);
    ac_dec st;
    st.p = bytes; st.bp = 0; st.n = nbytes;
    i16 lsb_ind[LC3O_MAX_N];
    int lsb_c = 0;
    const i16 rateFlag = nbbits > 160 + d->fs_idx * 160 ? 2 << 8 : 0;
    const int nt_half = d->yLen >> 1;
    i16 c = 0, t, a, b, a1, b1, esc_nb = 0, sym;
    int k, lev;
    memset(o, 0, sizeof(*o));
    ac_init(&st);
    int maxord = MAXLAG;
    if (d->dms == 25) maxord >>= 1;
    if (d->dms == 50) maxord >>= 1;
    for (int n = 0; n < si->nfilt; n++) {
        if (si->tns_order[n] > 0) {
            si->tns_order[n] = s_add(ac_decode_tns_order(&st, &rom_tns_order_cumfreq[lpc_weighting * 8]), 1);
            if (si->tns_order[n] > maxord) return 1;
            ac_update(&st, (u32)rom_tns_order_cumfreq[lpc_weighting * 8 + si->tns_order[n] - 1], (u32)rom_tns_order_freq[lpc_weighting * 8 + si->tns_order[n] - 1]);
            if (st.ber) return 1;
            for (k = 0; k < si->tns_order[n]; k++) {
                if (r->bp < st.bp) return 1;
                const i16 v = ac_decode17s(&st, &rom_tns_coef_cumfreq[k * 17]);
                o->tns_idx[MAXLAG * n + k] = v;
                ac_update(&st, (u32)rom_tns_coef_cumfreq[k * 17 + v], (u32)rom_tns_coef_freq[k * 17 + v]);
                if (st.ber) return 1;
            }
        }
    }
    if (st.ber > 0) return 1;

    for (k = 0; k < si->lastnz; k += 2) {
        t = s_add(c, rateFlag);
        if (k > nt_half) t = s_add(t, 1 << 8);
        int pki = (u8)rom_ari_lookup[t];
        sym = ac_decode17(&st, &rom_ari_cumfreq[pki * 17]);
        ac_update(&st, rom_ari_cumfreq[pki * 17 + sym], rom_ari_freq[pki * 17 + sym]);
        if (sym == 0) {
            a = 0; b = 0;
            c = s_add(s_shl(c & 0xf, 4), 1);
        } else if (sym < 16) {
            a = sym & 3; b = sym >> 2;
            c = s_add(s_shl(c & 0xf, 4), s_add(s_add(a, b), 1));
            if (a > 0 && rd_bit(r) != 0) a = s_neg(a);
            if (b > 0 && rd_bit(r) != 0) b = s_neg(b);
        } else {
            a = 0; b = 0;
            if (!si->lsbMode) { a = rd_bit(r); b = rd_bit(r); }
            pki = (u8)rom_ari_lookup[t + rom_ari_esc_nb[1]];
            sym = ac_decode17(&st, &rom_ari_cumfreq[pki * 17]);
            ac_update(&st, rom_ari_cumfreq[pki * 17 + sym], rom_ari_freq[pki * 17 + sym]);
            if (sym < 16) {
                a1 = sym & 3; b1 = sym >> 2;
                a = s_add(s_shl(a1, 1), a);
                b = s_add(s_shl(b1, 1), b);
                if (a > 0 && rd_bit(r) != 0) a = s_neg(a);
                if (b > 0 && rd_bit(r) != 0) b = s_neg(b);
                c = s_add(s_shl(c & 0xf, 4), s_add(s_shl(s_add(a1, b1), 1), 1));
            } else {
                a = s_add(s_shl(rd_bit(r), 1), a);
                b = s_add(s_shl(rd_bit(r), 1), b);
                for (lev = 2; lev < 14; lev++) {
                    esc_nb = s_min((i16)lev, 3);
                    pki = (u8)rom_ari_lookup[t + rom_ari_esc_nb[esc_nb]];
                    sym = ac_decode17(&st, &rom_ari_cumfreq[pki * 17]);
                    ac_update(&st, rom_ari_cumfreq[pki * 17 + sym], rom_ari_freq[pki * 17 + sym]);
                    if (sym < 16 || st.ber) break;
                    a = s_add(s_shl(rd_bit(r), (i16)lev), a);
                    b = s_add(s_shl(rd_bit(r), (i16)lev), b);
                }
                if (lev == 14 || st.ber) return 1;
                b1 = sym >> 2; a1 = sym & 3;
                a = s_add(s_shl(a1, (i16)lev), a);
                b = s_add(s_shl(b1, (i16)lev), b);
                if (a > 0 && rd_bit(r) != 0) a = s_neg(a);
                if (b > 0 && rd_bit(r) != 0) b = s_neg(b);
                c = s_add(s_shl(c & 0xf, 4), s_add(esc_nb, 12));
            }
            if (si->lsbMode) lsb_ind[lsb_c++] = (i16)k;
        }
        x[k] = a; x[k + 1] = b;
        if (s_sub(s_sub((i16)st.bp, r->bp), 3) > 0 || st.ber > 0 || r->oob) return 1;
    }
    for (; k < d->yLen; k++) x[k] = 0;

    const i16 nbits_side = s_sub(nbbits, s_add(s_shl(r->bp, 3), s_sub(s_norm(r->mask), 6)));
    const i16 extra_bits = s_sub(ul_norm(st.range), 6);
    const i16 nbits_ari = s_shl(s_sub((i16)st.bp, 3), 3);
    i16 n = s_sub(nbbits, s_add(nbits_ari, s_add(extra_bits, nbits_side)));
    if (n < 0) return 1;

    if (si->lsbMode == 0) {
        o->resBits = n;
        int cnt = 0;
        for (k = 0; k < d->yLen; k++)
            if (x[k] != 0) {
                if (n == 0) break;
                o->res[cnt++] = (u8)rd_bit(r);
                n = s_sub(n, 1);
            }
        o->resBits = s_sub(o->resBits, n);
    } else {
        o->resBits = 0;
        for (k = 0; k < lsb_c; k++) {
            a = x[lsb_ind[k]];
            if (n == 0) break;
            i16 tmp = rd_bit(r);
            n = s_sub(n, 1);
            if (tmp > 0) {
                if (a > 0) a = s_add(a, 1);
                if (a < 0) a = s_sub(a, 1);
                if (a == 0) {
                    if (n == 0) break;
                    a = 1;
                    if (rd_bit(r) != 0) a = s_neg(a);
                    n = s_sub(n, 1);
                }
            }
            x[lsb_ind[k]] = a;
            b = x[lsb_ind[k] + 1];
            if (n == 0) break;
            tmp = rd_bit(r);
            n = s_sub(n, 1);
            if (tmp > 0) {
                if (b > 0) b = s_add(b, 1);
                if (b < 0) b = s_sub(b, 1);
                if (b == 0) {
                    if (n == 0) break;
                    b = 1;
                    if (rd_bit(r) != 0) b = s_neg(b);
                    n = s_sub(n, 1);
                }
            }
            x[lsb_ind[k] + 1] = b;
        }
    }
    if (r->oob) return 1;
    i32 seed = 0;
    for (int i = 0; i < d->yLen; i++) seed = l_mac0(seed, s_abs(x[i]), (i16)i);
    o->nf_seed = lo16(seed);
    o->zero_frame = (si->lastnz == 2 && x[0] == 0 && x[1] == 0 && si->gg_idx == 0 && si->fac_ns_idx == 7) ? 1 : 0;
    return 0;
}

/* processAriDecoderScaling_fx, fx/ari_codec.c:1217-1256 */
static void ari_scaling(const i16 *x16, int n, i32 *x32, i16 *x_e)
{
    i16 x_max = 0, x_min = 0;
    for (int i = 0; i < n; i++) {
        if (x16[i] > 0) x_max = s_max(x_max, x16[i]);
        if (x16[i] < 0) x_min = s_min(x_min, x16[i]);
    }
    const i16 tmp = s_max(x_max, s_neg(x_min));
    i16 shift = s_norm(tmp);
    if (tmp == 0) shift = 15;
    for (int i = 0; i < n; i++) x32[i] = l_shl(l_dep_h(x16[i]), shift);
    *x_e = s_sub(15, shift);
}

/* processResidualDecoding_fx, fx/residual_decoding_fx.c:12-69 */
static void residual_decoding(i32 *x, i16 x_e, int L, const u8 *prm, i16 resQBits)
{
    const i16 s = s_sub(x_e, 1);
    const i32 fac_m = l_shr(0xC000000, s), fac_p = l_shr(0x14000000, s);
    i16 bits = 0;
    i32 tmp = 0;
    for (int i = 0; i < L; i++) {
        if (bits >= resQBits) break;
        if (x[i] != 0) {
            if (prm[bits] == 0) {
                if (x[i] > 0) tmp = l_sub(x[i], fac_m);
                if (x[i] < 0) tmp = l_sub(x[i], fac_p);
            } else {
                if (x[i] > 0) tmp = l_add(x[i], fac_p);
                if (x[i] < 0) tmp = l_add(x[i], fac_m);
            }
            x[i] = tmp;
            bits = s_add(bits, 1);
        }
    }
}

/* processNoiseFilling_fx, fx/noise_filling_fx.c:14-129 (spec_inv_idx = L_spec: no partial concealment) */
static void noise_filling(i32 *xq, i16 nfseed, i16 xq_e, i16 fac_ns_idx, i16 bw_idx, int dms)
{
    i16 ind[LC3O_MAX_N];
    int c = 0, N = rom_bw_cutoff_bin[bw_idx], w, start;
    if (dms == 25) { N >>= 2; w = 1; start = 6; }
    else if (dms == 50) { N >>= 1; w = 1; start = 12; }
    else { w = 3; start = 24; }
    i16 nzeros = (i16)(-2 * w - 1);
    int k;
    for (k = start - w; k < start + w; k++) {
        if (xq[k] != 0) nzeros = (i16)(-2 * w - 1);
        if (xq[k] == 0) nzeros = s_add(nzeros, 1);
    }
    for (k = start; k < N - w; k++) {
        if (xq[k + w] != 0) nzeros = (i16)(-2 * w - 1);
        if (xq[k + w] == 0) nzeros = s_add(nzeros, 1);
        if (nzeros >= 0) ind[c++] = (i16)k;
    }
    for (k = N - w; k < N; k++) {
        nzeros = s_add(nzeros, 1);
        if (nzeros >= 0) ind[c++] = (i16)k;
    }
    if (c > 0) {
        const i16 fac_ns = s_shl(s_sub(8, fac_ns_idx), 11);
        const i32 v = l_shr((i32)fac_ns, s_sub(xq_e, 16)), vn = l_neg(v);
        for (k = 0; k < c; k++) {
            nfseed = lo16(l_mac0(13849, nfseed, 31821));
            xq[ind[k]] = nfseed >= 0 ? v : vn;
        }
    }
}

/* processApplyGlobalGain_fx, fx/apply_global_gain_fx.c:14-44 */
static void apply_global_gain(i32 *x, i16 *x_e, int n, i16 gg_idx, i16 gg_off)
{
    const i32 t = l_shl(l_mult0(s_add(gg_idx, gg_off), 0x797D), 7);
    const i16 gg_e = s_add(lo16(l_shr(t, 25)), 1);
    const i16 gg = l_round(o_invlog2(t | (i32)0xFE000000));
    for (int i = 0; i < n; i++) x[i] = mpy32x16(x[i], gg);
    *x_e = s_add(*x_e, gg_e);
}

/* IIRLattice + processTnsDecoder_fx, fx/tns_decoder_fx.c:20-134 */
static i32 iir_lattice(int order, const i16 *rc, i32 *state, i32 x)
{
    x = l_sub(x, mpy32x16(state[order - 1], rc[order - 1]));
    for (int i = order - 2; i >= 0; i--) {
        x = l_sub(x, mpy32x16(state[i], rc[i]));
        state[i + 1] = l_add(state[i], mpy32x16(x, rc[i]));
    }
    state[0] = x;
    return x;
}
static void tns_decoder(const i16 *rc_idx, i32 *x, int xLen, const i16 *order, i16 *x_e, i16 bw_idx, int dms)
{
    i32 state[MAXLAG];
    i16 rc[MAXLAG], startfreq[2] = {12, 0};
    int numfilters = 1, bw_stop = rom_bw_cutoff_bin[bw_idx], stopfreq = 0, i, j;
    if (dms == 25) { startfreq[0] = 3; bw_stop >>= 2; }
    else if (dms == 50) { startfreq[0] = 6; bw_stop >>= 1; }
    if (bw_idx >= 3 && dms >= 50) { numfilters = 2; startfreq[1] = (i16)(bw_stop >> 1); }
    if (order[0] > 0 || (numfilters == 2 && order[1] > 0)) {
        int f = startfreq[0];
        if (numfilters == 2 && order[0] == 0) f = startfreq[1];
        const i16 s1 = o_headroom32(x, f, 0), s2 = o_headroom32(x + f, xLen - f, 0);
        const i16 s = s_min(s1, s_sub(s2, 7));
        *x_e = s_sub(*x_e, s);
        memset(state, 0, sizeof(state));
        for (i = 0; i < f; i++) x[i] = l_shl(x[i], s);
        for (j = 0; j < numfilters; j++)
            if (order[j] > 0) {
                for (i = 0; i < order[j]; i++) rc[i] = rom_tns_quant_pts[rc_idx[j * MAXLAG + i]];
                stopfreq = bw_stop;
                if (numfilters == 2 && j == 0) stopfreq = startfreq[1];
                for (i = startfreq[j]; i < stopfreq; i++) x[i] = iir_lattice(order[j], rc, state, l_shl(x[i], s));
            }
        for (i = stopfreq; i < xLen; i++) x[i] = l_shl(x[i], s);
    }
}

/* ---- SNS (fx/pvq_index_fx.c:90-120, :202-345; fx/pvq_dec_fx.c:15-52; fx/sns_quantize_scf_fx.c:442-570) */
static void a_bwd(u32 *a, int n_items)
{
    u32 a0 = 0;
    int i;
    for (i = 1; i <= n_items; i++) {
        const u32 a1 = a[i] - a0 - a[i - 1];
        a[i - 1] = a0;
        a0 = a1;
    }
    a[i - 1] = a0;
}
static void pvq_deidx(i16 *y, int k_val, int dim, i16 ls_ind, u32 index)
{
    u32 h[24];
    const uint32_t *offs = dim == 16 ? rom_mpvq_offs16 : (dim == 10 ? rom_mpvq_offs10 : rom_mpvq_offs6);
    const int kmax_tab = dim == 16 ? 12 : (dim == 10 ? 22 : 2), kp1 = k_val + 1;
    for (int i = 0; i <= kp1; i++) h[i] = offs[i];
    if (k_val != kmax_tab) h[kp1] >>= 1;
    const u32 size = 1u + h[kp1] + (h[k_val] >> 1);
    if (index >= size) index = 0;
    memset(y, 0, dim * sizeof(i16));
    i16 sign = ls_ind != 0 ? -1 : 1;
    if (k_val == 0) return;
    int k_max = k_val, k_acc = k_val, k_delta = 0;
    for (int pos = 0; pos < dim; pos++) {
        if (index != 0) {
            k_acc = k_max;
            while (index < h[k_acc]) k_acc--;
            index -= h[k_acc];
            k_delta = k_max - k_acc;
        } else {
            y[pos] = (i16)(sign < 0 ? -k_max : k_max);
            break;
        }
        if (k_delta != 0) {
            y[pos] = (i16)(sign < 0 ? -k_delta : k_delta);
            sign = (index & 1) ? -1 : 1;
            index >>= 1;
            k_max -= k_delta;
        }
        a_bwd(h, k_max + 1);
    }
}
static void sns_decode(const i32 *idx, i16 *scf_q)
{
    for (int i = 0; i < 8; i++) { scf_q[i] = rom_sns_cb_lf[idx[0] * 8 + i]; scf_q[8 + i] = rom_sns_cb_hf[idx[1] * 8 + i]; }
    const int submode = (int)idx[2];
    const i16 gain = rom_sns_gain[submode * 8 + idx[3]];
    i16 pulses[16], en1[16], adj[16];
    const i16 maxK = rom_sns_kval[submode * 2];
    if (submode >> 1) pvq_deidx(pulses, maxK, 16, (i16)idx[4], (u32)idx[5]);
    else {
        pvq_deidx(pulses, maxK, 10, (i16)idx[4], (u32)idx[5]);
        if ((submode & 1) == 0) pvq_deidx(pulses + 10, rom_sns_kval[submode * 2 + 1], 6, (i16)(idx[6] & 1), (u32)(idx[6] >> 1));
        else memset(pulses + 10, 0, 6 * sizeof(i16));
    }
    o_pvq_unit_norm(en1, pulses, maxK);
    o_idct16(en1, adj);
    for (int i = 0; i < 16; i++) scf_q[i] = s_add(scf_q[i], s_mult_r(gain, adj[i]));
}

/* ProcessingIMDCT, fx/imdct_fx.c:12-190 (regular operation) */
static void imdct(const lc3o_dec *d, i32 *y, i16 *y_e, i16 *mem, i16 *mem_e, i16 *x)
{
    const int N = d->N, z = d->la_zeroes, m = N >> 1, o = m - z, memLen = d->ola_len;
    const i16 *w = &rom_win[d->W_off];
    int i;
    for (i = d->yLen; i < N; i++) y[i] = 0;
    i16 y_s = o_headroom32(y, N, 1), mem_s;
    if (y_s < 32) {
        for (i = 0; i < N; i++) y[i] = l_shl(y[i], y_s);
        *y_e = s_sub(*y_e, y_s);
        o_dct4(y, y_e, N);
        y_s = s_sub(o_headroom32(y, N, 0), 1);
        *y_e = s_sub(*y_e, (i16)(y_s + 3));
        if (N <= 20) *y_e = s_add(*y_e, 2);
        else if (N <= 120) *y_e = s_add(*y_e, 1);
    } else *y_e = 0;
    mem_s = o_headroom16(mem, memLen, 1);
    if (mem_s < 16) { mem_s = s_sub(mem_s, 1); *mem_e = s_sub(*mem_e, mem_s); }
    else *mem_e = *y_e;
    const i16 s = s_sub(*mem_e, *y_e);
    if (s > 0) { y_s = s_sub(y_s, s); *y_e = s_add(*y_e, s); }
    else { mem_s = s_add(mem_s, s); *mem_e = s_sub(*mem_e, s); }
    mem_s = s_max(mem_s, -31);
    y_s = s_max(y_s, -31);
    for (i = 0; i < o; i++)
        x[i] = l_round(l_sub(l_shl(l_dep_h(mem[i]), mem_s), mpy32x16(l_shl(y[m + i + z], y_s), w[4 * m - 1 - i - z])));
    for (i = 0; i < m; i++)
        x[i + o] = l_round(l_add(l_shl(l_dep_h(mem[i + o]), mem_s), mpy32x16(l_shl(y[2 * m - 1 - i], y_s), w[3 * m - 1 - i])));
    for (i = 0; i < m; i++) x[3 * m - z + i] = l_round(l_neg(mpy32x16(l_shl(y[i], y_s), w[m - 1 - i])));
    for (i = 0; i < m; i++) x[3 * m - z - 1 - i] = l_round(l_neg(mpy32x16(l_shl(y[i], y_s), w[m + i])));
    memmove(mem, &x[N], memLen * sizeof(i16));
    *mem_e = *y_e;
}

/* ltpf_synth_filter, fx/ltpf_decoder_fx.c:268-366 */
static void ltpf_synth(i16 *synth_ltp, const i16 *synth, int length, i16 pitch_int, i16 pitch_fr, i16 gain, int sf, int fs_idx, int fade)
{
    const i16 *ifl = &rom_ltpf_dec_inter[(fs_idx * 4 + pitch_fr) * 12], *tfl = &rom_ltpf_dec_tilt[(fs_idx * 4 + sf) * 11];
    const int ilen = rom_ltpf_dec_inter_len[fs_idx], tlen = rom_ltpf_dec_tilt_len[fs_idx];
    const i16 *x0 = &synth_ltp[-pitch_int + rom_ltpf_dec_inter_shift[fs_idx]], *y0 = synth;
    i16 alpha = 0, step = 0;
    if (fade != 0) {
        if (fade < 0) alpha = 0x7FFF;
        if (length == 20) step = 1638;
        if (length == 40) step = 819;
        if (length == 60) step = 546;
        if (length == 80) step = 409;
        if (length == 120) step = 273;
        if (fade < 0) step = s_neg(step);
    }
    for (int j = 0; j < length; j++) {
        i32 s = l_mult(x0[0], ifl[0]);
        int l;
        for (l = 1; l < ilen; l++) s = l_mac(s, x0[-l], ifl[l]);
        for (l = 0; l < tlen; l++) s = l_msu(s, y0[-l], tfl[l]);
        const i16 i = s_msu_r(s, y0[-l], tfl[l]);
        i16 k = s_mult_r(gain, i);
        if (fade != 0) k = s_mult_r(k, alpha);
        synth_ltp[j] = s_add(synth[j], k);
        if (fade != 0) alpha = s_add(alpha, step);
        x0++; y0++;
    }
}

/* process_ltpf_decoder_fx, fx/ltpf_decoder_fx.c:22-262 with bfi = 0; x_in == y_out (in place), as Dec_LC3_Channel calls it */
static void ltpf_decoder(const lc3o_dec *d, int chn, i16 *x_e, i16 *xio, const i16 *ltpf_idx, i16 scale_fac_idx, lc3o_dec *dw)
{
    const int L = d->N, old_x_len = d->ltpf_x_len, old_y_len = d->ltpf_y_len, fs_idx = d->fs_idx;
    i16 *old_x = dw->ch[chn].ltpf_x, *old_y = dw->ch[chn].ltpf_y;
    i16 *old_e = &dw->ch[chn].ltpf_e, *old_pitch_int = &dw->ch[chn].pitch_int, *old_pitch_fr = &dw->ch[chn].pitch_fr;
    i16 *old_gain = &dw->ch[chn].gain, *mem_active = &dw->ch[chn].active, *old_sf = &dw->ch[chn].scale_idx;
    i16 ltpf_active = ltpf_idx[1], pitch_int = 0, pitch_fr = 0, gain = 0, z[128 + 16];
    int i;
    if (ltpf_idx[0] != 0) {
        const i16 pi = ltpf_idx[2];
        if (pi < 380) { pitch_int = (i16)((pi + 64) >> 2); pitch_fr = (i16)(pi - (pitch_int << 2) + 128); }
        else if (pi < 440) { pitch_int = (i16)((pi - 126) >> 1); pitch_fr = (i16)((pi << 1) - (pitch_int << 2) - 252); }
        else { pitch_int = (i16)(pi - 283); pitch_fr = 0; }
        i16 pitch = s_add(s_shl(pitch_int, 2), pitch_fr);
        pitch = s_mult_r(s_shl(pitch, 2), (i16)rom_ltpf_dec_pitch_scale[fs_idx]);
        pitch_int = s_shr(pitch, 2);
        pitch_fr = s_sub(pitch, s_shl(pitch_int, 2));
    }
    if (scale_fac_idx < 0) ltpf_active = 0;
    if (ltpf_active != 0) gain = rom_ltpf_dec_gain[scale_fac_idx];

    if (ltpf_active == 0 && *mem_active == 0) {
        const i16 s = s_sub(*old_e, *x_e);
        if (s > 0) {
            memmove(old_y, &old_y[L], (old_y_len - L) * sizeof(i16));
            if (s > 15) {
                memset(&old_y[old_y_len - L], 0, L * sizeof(i16));
                memset(old_x, 0, old_x_len * sizeof(i16));
            } else {
                for (i = 0; i < L; i++) old_y[i + old_y_len - L] = s_shr(xio[i], s);
                for (i = 0; i < old_x_len; i++) old_x[i] = s_shr(xio[i + L - old_x_len], s);
            }
        } else {
            if (s < -15) memset(old_y, 0, (old_y_len - L) * sizeof(i16));
            else for (i = 0; i < old_y_len - L; i++) old_y[i] = s_shl(old_y[i + L], s);
            memmove(&old_y[old_y_len - L], xio, L * sizeof(i16));
            memmove(old_x, &xio[L - old_x_len], old_x_len * sizeof(i16));
            *old_e = *x_e;
        }
        *old_pitch_int = pitch_int; *old_pitch_fr = pitch_fr; *old_gain = 0; *mem_active = 0;
    } else {
        i16 *x = old_x + old_x_len, *y = old_y + old_y_len;
        const int N4 = rom_ltpf_dec_overlap[fs_idx], N34 = L - N4, tl = rom_ltpf_dec_tilt_len[fs_idx];
        memmove(x, xio, L * sizeof(i16));
        const i16 s0 = s_sub(s_min(o_headroom16(old_x, old_x_len, 1), o_headroom16(old_y, old_y_len, 1)), 1);
        *old_e = s_sub(*old_e, s0);
        const i16 s1 = s_sub(o_headroom16(x, L, 0), 1);
        *x_e = s_sub(*x_e, s1);
        const i16 s = s_sub(*old_e, *x_e);
        if (s > 0) {
            o_scale_sig(x, L, s_sub(s1, s));
            o_scale_sig(old_x, old_x_len, s0);
            o_scale_sig(old_y, old_y_len, s0);
            *x_e = *old_e;
        } else {
            o_scale_sig(x, L, s1);
            o_scale_sig(old_x, old_x_len, s_add(s0, s));
            o_scale_sig(old_y, old_y_len, s_add(s0, s));
            *old_e = *x_e;
        }
        if (ltpf_active == 0) ltpf_synth(y, x, N4, *old_pitch_int, *old_pitch_fr, *old_gain, *old_sf, fs_idx, -1);
        else if (*mem_active == 0) ltpf_synth(y, x, N4, pitch_int, pitch_fr, gain, scale_fac_idx, fs_idx, 1);
        else if (pitch_int == *old_pitch_int && *old_pitch_fr == pitch_fr) ltpf_synth(y, x, N4, pitch_int, pitch_fr, gain, scale_fac_idx, fs_idx, 0);
        else {
            ltpf_synth(y, x, N4, *old_pitch_int, *old_pitch_fr, *old_gain, *old_sf, fs_idx, -1);
            memmove(z, y - tl, (N4 + tl) * sizeof(i16));
            ltpf_synth(y, z + tl, N4, pitch_int, pitch_fr, gain, scale_fac_idx, fs_idx, 1);
        }
        if (ltpf_active > 0) ltpf_synth(y + N4, x + N4, N34, pitch_int, pitch_fr, gain, scale_fac_idx, fs_idx, 0);
        else memmove(&y[N4], &x[N4], N34 * sizeof(i16));
        memmove(xio, y, L * sizeof(i16));
        memmove(old_x, &old_x[L], old_x_len * sizeof(i16));
        memmove(old_y, &old_y[L], old_y_len * sizeof(i16));
        *old_pitch_int = pitch_int; *old_pitch_fr = pitch_fr; *old_gain = gain; *mem_active = ltpf_active;
    }
    *old_sf = scale_fac_idx;
}

/* Dec_LC3_Channel, fx/dec_lc3.c:15-322 (bfi 0 in; a frame that does not decode: zero spectrum, returns 1) */
static int dec_channel(lc3o_dec *d, int chn, int bits_per_sample, const u8 *bytes, int nbytes, void *s_out)
{
    i16 xq[LC3O_MAX_N], x_fx[2 * LC3O_MAX_N], q_exp = 0;
    i32 q_d[LC3O_MAX_N];
    side_t si;
    ari_out ao;
    memset(&si, 0, sizeof(si));
    memset(q_d, 0, sizeof(q_d));
    /* update_dec_bitrate, fx/setup_dec_lc3.c:271-336 */
    const int total_bits = nbytes << 3;
    int lpcw = total_bits < 480, tb = total_bits, q = total_bits / (10 * (d->fs_idx + 1));
    if (q > 115) q = 115;
    const i16 gg_off = (i16)-(q + 105 + 5 * (d->fs_idx + 1));
    if (d->dms == 25) { lpcw = 0; tb = lo16(l_shr(l_mult0(19661, (i16)total_bits), 13)); }
    else if (d->dms == 50) { lpcw = total_bits < 240; tb = total_bits * 2 - 160; }
    const int tmp = 80 * d->fs_idx;
    const i16 sf_idx = tb < 320 + tmp ? 0 : (tb < 400 + tmp ? 1 : (tb < 480 + tmp ? 2 : (tb < 560 + tmp ? 3 : -1)));

    side_rd r;
    r.p = bytes; r.n = nbytes; r.bp = (i16)((total_bits - 1) >> 3); r.mask = 1; r.oob = 0;
    int bfi = nbytes <= 0 ? 1 : side_info(d, &r, &si);
    if (!bfi) bfi = ari_decode(d, bytes, nbytes, &r, &si, lpcw, xq, &ao);
    if (!bfi) {
        ari_scaling(xq, d->yLen, q_d, &q_exp);
        i16 scf_q[16], g[64], g_exp[64];
        sns_decode(si.scf, scf_q);
        residual_decoding(q_d, q_exp, d->yLen, ao.res, ao.resBits);
        if (ao.zero_frame == 0) noise_filling(q_d, ao.nf_seed, q_exp, si.fac_ns_idx, si.bw_idx, d->dms);
        apply_global_gain(q_d, &q_exp, d->yLen, si.gg_idx, gg_off);
        tns_decoder(ao.tns_idx, q_d, d->yLen, si.tns_order, &q_exp, si.bw_idx, d->dms);
        o_sns_interp_dir(scf_q, g, g_exp, d->nbands, 0);
        i16 emax = g_exp[0];                                   /* processScfScaling, fx/mdct_shaping_fx.c:37-67 */
        for (int i = 1; i < d->nbands; i++) emax = s_max(emax, g_exp[i]);
        for (int i = 0; i < d->nbands; i++) g_exp[i] = s_sub(g_exp[i], emax);
        q_exp = s_add(q_exp, emax);
        o_mdct_shape(q_d, g, g_exp, &rom_bands_offset[d->bands_off], d->nbands);
    } else {
        si.ltpf[0] = si.ltpf[1] = si.ltpf[2] = 0;
        q_exp = 0;
    }
    imdct(d, q_d, &q_exp, d->ch[chn].ola, &d->ch[chn].ola_e, x_fx);
    ltpf_decoder(d, chn, &q_exp, x_fx, si.ltpf, nbytes > 0 ? sf_idx : (i16)-1, d);
    /* output scaling, fx/dec_lc3.c:289-311 */
    if (bits_per_sample == 16) {
        const i16 scale = s_sub(15, q_exp);
        for (int i = 0; i < d->N; i++) ((int16_t *)s_out)[i] = l_round(l_shr(l_dep_h(x_fx[i]), scale));
    } else {
        const i16 scale = s_sub(s_sub(31 + 16, (i16)bits_per_sample), q_exp);
        const i32 offset = l_shr(32768, s_sub(16, scale));
        for (int i = 0; i < d->N; i++) ((int32_t *)s_out)[i] = l_shr(l_add(l_dep_h(x_fx[i]), offset), scale);
    }
    return bfi;
}

/* Dec_LC3, fx/dec_lc3.c:325-487 without error protection */
static int dec_frame(lc3o_dec *d, const uint8_t *in, int num_bytes, void **out, int bits)
{
    if (!d || !in || !out) return LC3O_NULL_ERROR;
    int bfi = 0;
    for (int ch = 0; ch < d->channels; ch++) {
        const int nb = num_bytes / d->channels + (ch < (num_bytes % d->channels));
        if (num_bytes > 0 && (nb < 20 || nb > 435)) return 7; /* LC3_NUMBYTES_ERROR */
        bfi |= dec_channel(d, ch, bits, in, nb, out[ch]);
        in += nb;
    }
    return bfi ? LC3O_DECODE_ERROR : LC3O_OK;
}
int lc3o_dec16(lc3o_dec *d, const uint8_t *in, int num_bytes, int16_t **out) { return dec_frame(d, in, num_bytes, (void **)out, 16); }
int lc3o_dec24(lc3o_dec *d, const uint8_t *in, int num_bytes, int32_t **out) { return dec_frame(d, in, num_bytes, (void **)out, 24); }
