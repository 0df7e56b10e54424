/* ORACLE (test infrastructure, not product code).
 * 12.8 kHz resampler + 50 Hz high-pass (fx/resamp12k8_fx.c:15-113), open-loop pitch analysis
 * (fx/olpa_fx.c:16-222) and the LTPF parameter coder (fx/ltpf_coder_fx.c:17-307). */
#include "lc3o_internal.h"

/* fx/resamp12k8_fx.c:15-113 */
void o_resamp12k8(const i16 *x, int x_len, o_pitch_state *st, int mem_in_len, i16 *y, int *y_len, int fs_idx, int dms)
{
    i16 buf[LC3O_MAX_N + 64];
    int upfac = rom_resamp_params[fs_idx * 4 + 0], delay = rom_resamp_params[fs_idx * 4 + 1];
    int off_int = rom_resamp_params[fs_idx * 4 + 2], off_frac = rom_resamp_params[fs_idx * 4 + 3];
    const i16 *filt = &rom_resamp_filt[fs_idx * 240];
    int len = 128 / 4 * (dms / 25);
    *y_len = len;

    memcpy(buf, st->r12k8_mem_in, mem_in_len * sizeof(i16));
    memcpy(buf + mem_in_len, x, x_len * sizeof(i16));
    memcpy(st->r12k8_mem_in, buf + x_len, mem_in_len * sizeof(i16));

    int ii = 1, fr = 0;
    for (int n = 0; n < len; n++) {
        const i16 *in = &buf[ii], *co = &filt[fr * delay * 2];
        i32 acc = l_mult0(in[0], co[0]);
        for (int m = 1; m < delay * 2; m++)
            if (co[m]) acc = l_mac0(acc, in[m], co[m]);
        y[n] = l_round(acc);
        ii += off_int;
        fr += off_frac;
        if (upfac - fr <= 0) { ii++; fr -= upfac; }
    }

    /* 2nd-order high-pass, 32-bit state */
    i32 m0 = st->r12k8_mem_50[0], m1 = st->r12k8_mem_50[1];
    for (int n = 0; n < len; n++) {
        i32 out = l_mac0(m0, rom_hp50_num[0], y[n]);
        i32 t = l_mac0(mpy32x16(out, rom_hp50_den[0]), rom_hp50_num[1], y[n]);
        m0 = l_add(m1, l_shl(t, 1));
        m1 = l_mac0(mpy32x16(out, rom_hp50_den[1]), rom_hp50_num[2], y[n]);
        y[n] = l_round(out);
    }
    st->r12k8_mem_50[0] = m0;
    st->r12k8_mem_50[1] = m1;

    /* 24-sample output delay line; y gets len+1 samples */
    i16 ob[24 + 128];
    memcpy(ob, st->r12k8_mem_out, 24 * sizeof(i16));
    memcpy(ob + 24, y, len * sizeof(i16));
    memcpy(y, ob, (len + 1) * sizeof(i16));
    memcpy(st->r12k8_mem_out, ob + len, 24 * sizeof(i16));
}

/* normalised correlation <a,b>/sqrt(<a,a><b,b>) in Q15 (fx/olpa_fx.c:126-155, ltpf_coder_fx.c:108-141) */
static i16 normcorr_q15(i32 sum0, i32 sum1, i32 sum2, int sum1_prenorm, i16 sc1_in)
{
    i16 sc1 = sc1_in, sc2;
    if (!sum1_prenorm) { sc1 = l_norm(sum1); sum1 = l_shl(sum1, sc1); }
    sc2 = l_norm(sum2);
    sum2 = l_shl(sum2, sc2);
    i32 prod = mpy32x32(sum1, sum2);
    i16 sh = l_norm(prod);
    prod = l_shl(prod, sh);
    i16 pe = s_sub(62, s_add(s_add(sc1, sc2), sh));
    i32 inv = o_isqrt32(prod, &pe);
    i16 sc0 = l_norm(sum0);
    sum0 = l_shl(sum0, sc0);
    prod = mpy32x32(sum0, inv);
    pe = s_add(s_sub(31, sc0), pe);
    if (prod == 0 || s_sub(l_norm(prod), pe) >= 0) return s_max(0, l_round(l_shl(prod, pe)));
    return 32767;
}

static i16 olpa_corr_at(const i16 *s, int lag, int n)
{
    i32 s0 = l_mult0(s[0], s[-lag]);
    i32 s1 = l_mac0(1, s[-lag], s[-lag]);
    i32 s2 = l_mac0(1, s[0], s[0]);
    for (int m = 1; m < n; m++) {
        s0 = l_mac0(s0, s[m], s[m - lag]);
        s1 = l_mac0(s1, s[m - lag], s[m - lag]);
        s2 = l_mac0(s2, s[m], s[m]);
    }
    return normcorr_q15(s0, s1, s2, 0, 0);
}

/* fx/olpa_fx.c:16-222 */
void o_olpa(o_pitch_state *st, i16 *pitch, const i16 *s12, int len, i16 *normcorr, i16 s12k8_exp, int dms)
{
    i32 ac[98];
    int mem_len = 114, len2 = len >> 1, acflen = len2;
    if (dms == 25) { mem_len += 16; acflen += 16; }
    i16 *mem = st->olpa_mem_s6k4, *m12 = st->olpa_mem_s12k8;
    i16 *s6 = mem + mem_len, *w = s6;

    /* 2:1 decimation with a 5-tap low-pass */
    i32 sum = l_mac(l_mac(l_mult(m12[0], 4053), m12[1], 7712), m12[2], 9239);
    *w++ = l_round(l_mac(l_mac(sum, s12[0], 7712), s12[1], 4053));
    sum = l_mac(l_mac(l_mult(m12[2], 4053), s12[0], 7712), s12[1], 9239);
    *w++ = l_round(l_mac(l_mac(sum, s12[2], 7712), s12[3], 4053));
    for (int n = 5; n < len; n += 2) {
        sum = l_mac(l_mac(l_mult(s12[n - 4], 4053), s12[n - 3], 7712), s12[n - 2], 9239);
        *w++ = l_round(l_mac(l_mac(sum, s12[n - 1], 7712), s12[n], 4053));
    }
    m12[0] = s12[len - 3]; m12[1] = s12[len - 2]; m12[2] = s12[len - 1];

    /* bring history and new samples to a common block exponent */
    i16 sc0 = s_sub(o_headroom16(mem, mem_len, 16), 3);
    st->olpa_mem_s6k4_exp = s_sub(st->olpa_mem_s6k4_exp, sc0);
    i16 sc1 = s_sub(o_headroom16(s6, len2, 16), 3);
    i16 s6_exp = s_sub(s12k8_exp, sc1);
    i16 sc2 = s_sub(st->olpa_mem_s6k4_exp, s6_exp);
    i16 shift;
    if (sc2 > 0) {
        o_scale_sig(s6, len2, s_sub(sc1, sc2));
        shift = sc0;
    } else {
        o_scale_sig(s6, len2, sc1);
        shift = s_add(sc0, sc2);
        st->olpa_mem_s6k4_exp = s6_exp;
    }
    if (dms == 25) s6 -= 16;
    o_scale_sig(mem, mem_len, shift);

    for (int n = 17; n <= 114; n++) {
        sum = l_mult0(s6[0], s6[-n]);
        for (int m = 1; m < acflen; m++) sum = l_mac0(sum, s6[m], s6[m - n]);
        ac[n - 17] = sum;
    }

    /* weighted maximum; a later lag wins only if strictly greater */
    i32 mx = mpy32x16(ac[0], rom_olpa_weight[0]);
    i16 p = 17;
    for (int n = 18; n <= 114; n++) {
        i32 t = mpy32x16(ac[n - 17], rom_olpa_weight[n - 17]);
        if (l_sub(t, mx) > 0) p = (i16)n;
        mx = l_max(t, mx);
    }
    *normcorr = olpa_corr_at(s6, p, acflen);

    /* second candidate around the previous pitch */
    i16 lo = s_max(17, s_sub(st->olpa_mem_pitch, 4)), hi = s_min(114, s_add(st->olpa_mem_pitch, 4));
    mx = ac[lo - 17];
    i16 p2 = lo;
    for (int n = lo + 1; n <= hi; n++) {
        if (l_sub(ac[n - 17], mx) > 0) p2 = (i16)n;
        mx = l_max(ac[n - 17], mx);
    }
    if (p != p2) {
        i16 nc2 = olpa_corr_at(s6, p2, acflen);
        if (s_sub(nc2, s_mult_r(*normcorr, 27853)) > 0) { p = p2; *normcorr = nc2; }
    }
    st->olpa_mem_pitch = p;
    memmove(mem, mem + len2, mem_len * sizeof(i16));
    *pitch = s_shl(p, 1);
}

/* fx/ltpf_coder_fx.c:17-307 */
void o_ltpf(o_pitch_state *st, i16 *bits, i16 ol_pitch, i16 ltpf_enable, int mem_in_len, i16 param[3],
            const i16 *xin, int len, i16 ol_normcorr, i16 xin_exp, int dms)
{
    i16 ac[17], cur[128], pred[128];
    i16 *mem = st->ltpf_mem_in, *x = mem + mem_in_len;
    i16 ltpf_active = 0, norm_corr = 0, ltpf_pitch;

    memcpy(x, xin, (len + 1) * sizeof(i16));

    i16 sc0 = s_sub(o_headroom16(mem, mem_in_len, 16), 3);
    st->ltpf_mem_in_exp = s_sub(st->ltpf_mem_in_exp, sc0);
    i16 sc1 = s_sub(o_headroom16(x, len + 1, 16), 3);
    i16 x_exp = s_sub(xin_exp, sc1);
    i16 sc2 = s_sub(st->ltpf_mem_in_exp, x_exp);
    if (sc2 > 0) {
        o_scale_sig(x, len + 1, s_sub(sc1, sc2));
        o_scale_sig(mem, mem_in_len, sc0);
    } else {
        o_scale_sig(x, len + 1, sc1);
        o_scale_sig(mem, mem_in_len, s_add(sc0, sc2));
        st->ltpf_mem_in_exp = x_exp;
    }

    if (s_sub(ol_normcorr, 19660) > 0) {
        i16 min_p = s_max(s_sub(ol_pitch, 4), 32), max_p = s_min(s_add(ol_pitch, 4), 228);
        i16 ac_lo = s_sub(min_p, 4), ac_hi = s_add(max_p, 4);
        int acflen = len;
        if (dms == 25) { acflen = len * 2; x -= len; }

        i32 sum1 = l_mac0(1, x[0], x[0]), sum2 = l_mac0(1, x[-ac_lo], x[-ac_lo]);
        for (int m = 1; m < acflen; m++) {
            sum1 = l_mac0(sum1, x[m], x[m]);
            sum2 = l_mac0(sum2, x[m - ac_lo], x[m - ac_lo]);
        }
        i16 n1 = l_norm(sum1);
        sum1 = l_shl(sum1, n1);

        for (int n = ac_lo; n <= ac_hi; n++) {
            i32 sum0 = l_mac0(0, x[0], x[-n]);
            for (int m = 1; m < acflen; m++) sum0 = l_mac0(sum0, x[m], x[m - n]);
            if (n > ac_lo) { /* slide the lagged-energy window by one sample */
                sum2 = l_msu0(sum2, x[acflen - 1 - (n - 1)], x[acflen - 1 - (n - 1)]);
                sum2 = l_mac0(sum2, x[-n], x[-n]);
            }
            ac[n - ac_lo] = normcorr_q15(sum0, sum1, sum2, 1, n1);
        }

        i16 ac_max = ac[min_p - ac_lo], pitch = min_p;
        for (int n = min_p + 1; n <= max_p; n++) {
            if (s_sub(ac[n - ac_lo], ac_max) > 0) pitch = (i16)n;
            ac_max = s_max(ac_max, ac[n - ac_lo]);
        }
        i16 p_int = pitch, p_fr = 0, p_idx = s_add(p_int, 283);

        if (pitch < 157) { /* fractional refinement */
            int res = pitch < 127 ? 1 : 2, fr0 = pitch < 127 ? -3 : -2;
            if (pitch == min_p) fr0 = 0;
            i32 best = I32_MIN_;
            for (int fr = fr0; fr < 4; fr += res) {
                const i16 *h = &rom_ltpf_ac_interp[(fr + 3) * 9];
                i32 s = l_mult0(ac[p_int - ac_lo - 4], h[0]);
                for (int k = 1; k < 9; k++) s = l_mac0(s, ac[p_int - ac_lo - 4 + k], h[k]);
                if (l_sub(s, best) > 0) p_fr = (i16)fr;
                best = l_max(best, s);
            }
            if (p_fr < 0) { p_int = s_sub(p_int, 1); p_fr = s_add(p_fr, 4); }
            if (p_int >= 127) p_idx = s_add(s_add(s_shl(p_int, 1), s_shr(p_fr, 1)), 126);
            else p_idx = s_sub(s_add(s_shl(p_int, 2), p_fr), 128);
        }
        ltpf_pitch = s_add(s_shl(p_int, 2), p_fr);

        const i16 *f0 = &rom_ltpf_inter_filt[0], *fp = &rom_ltpf_inter_filt[p_fr * 12];
        for (int n = 0; n < acflen; n++) {
            i32 s = l_mult(x[n + 1], f0[0]);
            s = l_mac(s, x[n], f0[1]);
            cur[n] = s_mac_r(s, x[n - 1], f0[2]);
            s = l_mult(x[n - p_int + 1], fp[0]);
            s = l_mac(s, x[n - p_int], fp[1]);
            s = l_mac(s, x[n - p_int - 1], fp[2]);
            pred[n] = s_mac_r(s, x[n - p_int - 2], fp[3]);
        }
        i32 sum0 = l_mult0(cur[0], pred[0]);
        sum1 = l_mac0(1, pred[0], pred[0]);
        sum2 = l_mac0(1, cur[0], cur[0]);
        for (int m = 1; m < acflen; m++) {
            sum0 = l_mac0(sum0, cur[m], pred[m]);
            sum1 = l_mac0(sum1, pred[m], pred[m]);
            sum2 = l_mac0(sum2, cur[m], cur[m]);
        }
        norm_corr = normcorr_q15(sum0, sum1, sum2, 0, 0);
        if (norm_corr < 0) norm_corr = 0;

        if (ltpf_enable == 1) {
            i16 on = st->ltpf_mem_ltpf_on, mnc = st->ltpf_mem_normcorr;
            if ((on == 0 && s_sub(mnc, 30802) > 0 && s_sub(norm_corr, 30802) > 0 &&
                 (dms == 100 || s_sub(st->ltpf_mem_mem_normcorr, 30802) > 0)) ||
                (on == 1 && s_sub(norm_corr, 29491) > 0) ||
                (on == 1 && s_sub(s_abs(s_sub(ltpf_pitch, st->ltpf_mem_pitch)), 8) < 0 &&
                 s_add(s_sub(norm_corr, mnc), 3277) > 0 && s_sub(norm_corr, 27525) > 0))
                ltpf_active = 1;
        }
        param[0] = 1; param[1] = ltpf_active; param[2] = p_idx;
        *bits = 11;
    } else {
        norm_corr = ol_normcorr;
        param[0] = param[1] = param[2] = 0;
        *bits = 1;
        ltpf_pitch = 0;
    }

    for (int n = 0; n < mem_in_len; n++) mem[n] = mem[n + len];
    st->ltpf_mem_mem_normcorr = st->ltpf_mem_normcorr;
    st->ltpf_mem_normcorr = norm_corr;
    st->ltpf_mem_ltpf_on = ltpf_active;
    st->ltpf_mem_pitch = ltpf_pitch;
}
