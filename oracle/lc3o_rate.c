/* ORACLE (test infrastructure, not product code).
 * Rate loop: global-gain estimation (fx/estimate_global_gain_fx.c:15-210), spectral quantiser with
 * its context bit counter (fx/quantize_spec_fx.c:15-553, fx/ari_codec.c:1269-1308), gain adjustment
 * (fx/adjust_global_gain_fx.c:14-83), residual bits (fx/residual_coding_fx.c:14-75) and the
 * noise-level estimate (fx/noise_factor_fx.c:14-213). */
#include "lc3o_internal.h"

/* gain index -> linear gain mantissa/exponent (estimate_global_gain_fx.c:200-204, adjust_...:72-75) */
static void gain_from_index(i16 gi, i16 *gain, i16 *gain_e)
{
    i32 t = l_shl(l_mult0(gi, 0x797D), 7);
    *gain_e = s_add(lo16(l_shr(t, 25)), 1);
    *gain = l_round(o_invlog2(t | (i32)0xFE000000));
}

/* fx/estimate_global_gain_fx.c:15-210 */
void o_est_gain(const i32 *x, i16 x_e, int lg, i16 nbitsSQ, i16 *gain, i16 *gain_e, i16 *qgain, i16 *qgain_min,
                i16 qgain_off, i32 *tbits_off, i16 *old_tbits, i16 old_specbits)
{
    i32 en[LC3O_MAX_N / 4];
    int lg4 = lg >> 2;

    /* rate-loop feedback: low-pass of (last target - last consumed), clamped to +-40 bits */
    if (*old_tbits < 0) {
        *tbits_off = 0;
    } else {
        i32 t = l_add(*tbits_off, l_dep_h(s_sub(*old_tbits, old_specbits)));
        t = l_min(40 << 16, l_max(-(40 << 16), t));
        *tbits_off = l_add(mpy32x16(*tbits_off, 26214), mpy32x16(t, 6554));
    }
    *old_tbits = nbitsSQ;
    nbitsSQ = s_add(nbitsSQ, l_round(*tbits_off));

    i32 x_max = 0;
    for (int i = 0; i < lg4; i++, x += 4) { /* log-energy of each quadruple, (28/20)*(7+10log10) Q16 */
        i32 m = l_max(l_max(l_abs(x[0]), l_abs(x[1])), l_max(l_abs(x[2]), l_abs(x[3])));
        x_max = l_max(x_max, m);
        i16 s = m != 0 ? l_norm(m) : 31;
        s = s_sub(s, 2);
        i32 e = 1;
        for (int k = 0; k < 4; k++) { i16 t = l_round(l_shl(x[k], s)); e = l_mac(e, t, t); }
        s = s_shl(s_sub(x_e, s), 1);
        if (e == 1 && s < 0) s = 0;
        i32 t = l_add(o_log2(e), l_shl((i32)s, 25));
        en[i] = l_add(l_shr(mpy32x16(t, 0x436E), 6), 0x9CCCD);
    }

    if (x_max == 0) {
        *qgain_min = qgain_off;
        *qgain = 0;
        *old_tbits = -1;
    } else {
        x_max = l_add(o_log2(x_max), l_shl((i32)x_e, 25));
        x_max = l_sub(mpy32x32(x_max, 0x436E439A), 0xFCDD6B4);
        *qgain_min = lo16(l_shr(l_add(x_max, (1 << 21) + (1 << 11)), 21));
        *qgain_min = s_max(qgain_off, s_min(s_add(255, qgain_off), *qgain_min));

        i32 offset = l_dep_h(s_add(255, qgain_off));
        i32 target = l_shl(l_mult(0x7D71, nbitsSQ), 1), fac = 0x1000000;
        for (int it = 0; it < 8; it++) { /* bisection on the gain index */
            fac = l_shr(fac, 1);
            offset = l_sub(offset, fac);
            i32 est = 0;
            int all_zero = 1;
            for (int i = lg4 - 1; i >= 0; i--) {
                i32 t = l_sub(en[i], offset);
                if (l_sub(t, 0x9CCCD) < 0) {
                    if (!all_zero) est = l_add(est, 0x3C7AE);
                } else {
                    est = l_add(est, t);
                    all_zero = 0;
                    i32 d2 = l_sub(t, 0x460000);
                    if (d2 >= 0) est = l_add(est, d2);
                }
            }
            if (l_sub(est, target) > 0 && !all_zero) offset = l_add(offset, fac);
        }
        i16 g = hi16(offset);
        if (s_sub(g, *qgain_min) < 0) *old_tbits = -1;
        *qgain = s_sub(s_max(*qgain_min, g), qgain_off);
    }
    gain_from_index(s_add(*qgain, qgain_off), gain, gain_e);
}

/* fx/ari_codec.c:1269-1308 */
static i16 last_nz_pair(const i16 *x, int n)
{
    int lobs[4] = {4, n >> 1, (n >> 1) + (n >> 2), (n >> 1) + (n >> 2) + (n >> 3)};
    i16 last = 0;
    int i = n;
    for (int stage = 3; stage >= 0; stage--) {
        for (; i >= lobs[stage]; i -= 2)
            if (x[i - 2] != 0 || x[i - 1] != 0) last = s_max(last, (i16)i);
        if (last > 0) break;
    }
    return s_max(last, 2);
}

static inline i32 ari_bits(int t, int sym) { return rom_ari_bits[rom_ari_lookup[t] * 17 + sym]; }

/* fx/quantize_spec_fx.c:15-553.  The reference unrolls three copies of the tuple loop for
 * mode -1 (first pass), 0 (second pass, truncation tracking) and 1 (plus LSB-plane accounting);
 * they are folded into one loop here. */
void o_quantize(const i32 *x, i16 x_e, i16 gain, i16 gain_e, i16 *xq, int nt, i16 target, i16 totalBits,
                i16 *nBits, i16 *nBits2, int fs_idx, i16 *lastnz_out, i16 *cd, i16 *lsbMode, int mode)
{
    gain = o_inv16(gain, &gain_e);
    i16 s = s_sub(s_add(x_e, gain_e), 15);
    for (int i = 0; i < nt; i++) { /* |x|/gain + 0.375, truncated, sign restored */
        i32 v = l_shl(mpy32x16(l_abs(x[i]), gain), s);
        i16 q = s_mac_r(v, -4096, 1);
        xq[i] = x[i] < 0 ? s_neg(q) : q;
    }

    i16 rateFlag = s_sub(totalBits, s_add(160, s_imult((i16)fs_idx, 160))) > 0 ? 2 << 8 : 0;
    if (mode == 0 && s_sub(totalBits, s_add(480, s_imult((i16)fs_idx, 160))) >= 0) mode = 1;

    int nt_half = nt >> 1;
    i16 c = 0, nbits_lsb = 0;
    i32 target32 = l_shl((i32)target, 11), nb = l_neg(target32), nb2 = 0;
    i16 lastnz = last_nz_pair(xq, nt), lastnz2 = mode >= 0 ? 2 : lastnz;

    for (int k = 0; k < lastnz; k += 2, cd += 3) {
        i16 t = s_add(c, rateFlag);
        if (k > nt_half) t = s_add(t, 1 << 8);
        cd[0] = t;
        i16 a1 = s_abs(xq[k]), b1 = s_abs(xq[k + 1]), mx = s_max(a1, b1);
        if (mx == 0) {
            cd[1] = -1; cd[2] = 0;
            nb = l_add(nb, ari_bits(t, 0));
            c = s_add(s_shl((i16)(c & 0xf), 4), 1);
            continue; /* an all-zero tuple never moves the truncation point */
        }
        i16 lev, am, bm;
        if (mx < 4) {
            lev = 0; am = a1; bm = b1;
        } else if (mx < 8) {
            lev = 1; am = s_shr(a1, 1); bm = s_shr(b1, 1);
            nb = l_add(nb, ari_bits(t, 16));
            if (mode <= 0) nb = l_add(nb, 2 << 11);
        } else {
            lev = s_sub(13, s_norm(mx));
            for (int l = 0; l < lev; l++) nb = l_add(nb, ari_bits(t + rom_ari_esc_nb[l < 3 ? l : 3], 16));
            nb = l_add(nb, l_shl((i32)(mode <= 0 ? lev : s_sub(lev, 1)), 12));
            am = s_shr(a1, lev); bm = s_shr(b1, lev);
        }
        cd[1] = lev;
        i16 msb = s_add(am, s_shl(bm, 2));
        cd[2] = msb;
        nb = l_add(nb, ari_bits(t + (lev ? rom_ari_esc_nb[lev < 3 ? lev : 3] : 0), msb));
        if (mode > 0 && lev > 0) { /* sign bits only for values that survive dropping the LSB */
            if (s_shr(a1, 1) != 0) nb = l_add(nb, 1 << 11);
            if (s_shr(b1, 1) != 0) nb = l_add(nb, 1 << 11);
            nbits_lsb = s_add(nbits_lsb, 2);
            if (a1 == 1) nbits_lsb = s_add(nbits_lsb, 1);
            if (b1 == 1) nbits_lsb = s_add(nbits_lsb, 1);
        } else {
            if (a1 != 0) nb = l_add(nb, 1 << 11);
            if (b1 != 0) nb = l_add(nb, 1 << 11);
        }
        if (mode >= 0 && nb <= 0) { lastnz2 = (i16)(k + 2); nb2 = nb; }
        if (lev == 0) c = s_add(s_shl((i16)(c & 0xf), 4), s_add(s_add(a1, b1), 1));
        else if (lev == 1) c = s_add(s_shl((i16)(c & 0xf), 4), s_add(s_shl(s_add(am, bm), 1), 1));
        else c = s_add(s_shl((i16)(c & 0xf), 4), s_add(12, s_min(lev, 3)));
    }

    nb = l_add(nb, target32);
    i16 tot = s_add(lo16(l_shr(l_sub(nb, 1), 11)), 1);
    if (mode > 0) tot = s_add(tot, nbits_lsb);
    if (nBits) *nBits = tot;
    if (mode >= 0) {
        nb2 = l_add(nb2, target32);
        *nBits2 = s_add(lo16(l_shr(l_sub(nb2, 1), 11)), 1);
    } else {
        *nBits2 = tot;
    }
    if (mode > 0) *nBits2 = s_add(*nBits2, nbits_lsb);
    *lastnz_out = lastnz2;
    if (lastnz > lastnz2) memset(&xq[lastnz2], 0, (lastnz - lastnz2) * sizeof(i16));
    *lsbMode = (mode > 0 && s_sub(tot, target) > 0) ? 1 : 0;
}

/* fx/adjust_global_gain_fx.c:14-83 */
void o_adjust_gain(i16 *gg, i16 gg_min, i16 gg_off, i16 *gain, i16 *gain_e, i16 target, i16 nBits, i16 *change, int fs_idx)
{
    const i16 *T = rom_adjust_gain;
    i16 delta;
    if (nBits < T[0 * 5 + fs_idx]) delta = s_mult_r(s_add(nBits, 48), 2048);
    else if (nBits < T[1 * 5 + fs_idx]) delta = s_mult_r(s_add(nBits, T[4 * 5 + fs_idx]), T[3 * 5 + fs_idx]);
    else if (nBits < T[2 * 5 + fs_idx]) delta = s_mult_r(nBits, 683);
    else delta = s_mult_r(T[2 * 5 + fs_idx], 683);
    i16 delta2 = s_add(delta, 2);

    *change = 0;
    if (*gg == 255 && nBits > target) *change = 1;
    if ((*gg < 255 && nBits > target) || (*gg > 0 && nBits < s_sub(target, delta2))) {
        if (nBits < s_sub(target, delta2)) *gg = s_sub(*gg, 1);
        else if (*gg == 254 || nBits < s_add(target, delta)) *gg = s_add(*gg, 1);
        else *gg = s_add(*gg, 2);
        *gg = s_max(*gg, s_sub(gg_min, gg_off));
        gain_from_index(s_add(*gg, gg_off), gain, gain_e);
        *change = 1;
    }
}

/* fx/residual_coding_fx.c:14-75 */
void o_residual(i16 x_e, const i32 *x, const i16 *xq, i16 gain, i16 gain_e, int L, i16 target, i16 nBits, uint8_t *res, i16 *nres)
{
    i16 n = 0, m = s_add(s_sub(target, nBits), 4);
    if (m > L) m = (i16)L;
    i16 s = s_sub(s_add(15, gain_e), x_e);
    for (int i = 0; i < L; i++) {
        if (xq[i] == 0) continue;
        res[n] = l_sub(x[i], l_shl(l_mult(xq[i], gain), s)) >= 0;
        n = s_add(n, 1);
        if (n == m) break;
    }
    *nres = n;
}

/* mean |x| over the selected bins / gain, in Q? per noise_factor_fx.c:186-204 */
static i16 level_over_gain(i32 sum, i16 cnt, i16 gg, i16 gg_e, i16 x_e, i16 norm_of)
{
    i16 s1, s2;
    i16 f = o_div3216(sum, cnt, &s1);
    f = o_div1616(f, gg, &s2);
    i16 s3 = s_sub(15, s_add(x_e, s_add(s1, s_sub(s2, gg_e))));
    s2 = s_norm(norm_of == I16_MIN_ ? f : norm_of); /* the reference normalises fac_unq1 in both halves */
    if (f != 0 && s_add(s3, s2) < 0) return I16_MAX_;
    return s_shr_r(f, s_min(s3, 15));
}

/* fx/noise_factor_fx.c:14-213 */
i16 o_noise_factor(i16 x_e, const i32 *x, const i16 *xq, i16 gg, i16 gg_e, int bw, int dms, i16 target_bytes)
{
    i16 ind[LC3O_MAX_N];
    int width = dms == 100 ? 3 : 1, start = dms == 100 ? 24 : (dms == 50 ? 12 : 6);
    int N = rom_bw_cutoff_bin[bw] >> (dms == 100 ? 0 : (dms == 50 ? 1 : 2));
    int c = 0, k;
    i16 nz = (i16)(-2 * width - 1);
    for (k = start - width; k < start + width; k++) nz = xq[k] != 0 ? (i16)(-2 * width - 1) : s_add(nz, 1);
    for (k = start; k < N - width; k++) {
        nz = xq[k + width] != 0 ? (i16)(-2 * width - 1) : s_add(nz, 1);
        if (nz >= 0) ind[c++] = (i16)k;
    }
    for (k = N - width; k < N; k++) {
        nz = s_add(nz, 1);
        if (nz >= 0) ind[c++] = (i16)k;
    }

    i16 fac = 0;
    if (c != 0) {
        if (target_bytes <= 20 && dms == 100) { /* split low/high around the mean index, take the min */
            i32 isum = ind[0];
            for (k = 1; k < c; k++) isum = l_add(isum, ind[k]);
            i16 s2;
            i16 mean = o_div3216(isum, (i16)c, &s2);
            mean = s_shl(mean, (i16)(s2 + 1));
            i32 sum = 0;
            for (k = 0; ind[k] <= mean; k++) sum = l_add(sum, l_abs(x[ind[k]]));
            i16 f1 = level_over_gain(sum, (i16)k, gg, gg_e, x_e, I16_MIN_);
            i16 cnt2 = (i16)(c - k);
            sum = 0;
            for (; k < c; k++) sum = l_add(sum, l_abs(x[ind[k]]));
            i16 f2 = level_over_gain(sum, cnt2, gg, gg_e, x_e, f1);
            fac = s_min(f1, f2);
        } else {
            i32 sum = l_abs(x[ind[0]]);
            for (k = 1; k < c; k++) sum = l_add(sum, l_abs(x[ind[k]]));
            fac = level_over_gain(sum, (i16)c, gg, gg_e, x_e, I16_MIN_);
        }
    }
    i16 idx = l_round(l_sub(0x80000, l_mult(fac, 16)));
    if (idx > 7) idx = 7;
    if (idx < 0) idx = 0;
    return idx;
}
