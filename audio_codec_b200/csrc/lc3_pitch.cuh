// Per-stream time-domain chain, one lane per stream: 12.8 kHz resampler + 50 Hz high-pass
// (fx/resamp12k8_fx.c:15-113), open-loop pitch analysis (fx/olpa_fx.c:16-222), LTPF parameter
// coder (fx/ltpf_coder_fx.c:17-307) and attack detector (fx/attack_detector_fx.c:13-119).
// This is the part of the encoder whose state makes frame n+1 depend on frame n, so the lane walks
// its stream's frames in order with the state held in PitchState.
#pragma once
#include "lc3_math.cuh"

namespace lc3k {

struct PitchState {
    i32 r12k8_mem_50[2];
    i16 r12k8_mem_out[24];
    i16 olpa_mem_s12k8[3];
    i16 olpa_buf[130 + 64];        // history (114 | 130) followed by the new 6.4 kHz samples
    i16 ltpf_buf[264 + 128 + 1];   // history (232 | 264) followed by the new 12.8 kHz samples
    i32 olpa_exp, olpa_pitch;
    i32 ltpf_exp, ltpf_normcorr, ltpf_mem_normcorr, ltpf_on, ltpf_pitch;
};
struct AttackState {
    i32 filter_mem[2];
    i32 detected, position, scaling, acc_energy;
};

// resampler + high-pass + 24-sample delay.  XS(i): scaled input sample, i in [-mem_in_len, N).
// out: len+1 samples of the delayed 12.8 kHz signal.
template <class XS> LC3_HD void resamp12k8(XS xs, int mem_in_len, PitchState &st, i16 *out, int len, int fs_idx)
{
    const int upfac = rom_resamp_params[fs_idx * 4 + 0], delay2 = 2 * rom_resamp_params[fs_idx * 4 + 1];
    const int off_int = rom_resamp_params[fs_idx * 4 + 2], off_frac = rom_resamp_params[fs_idx * 4 + 3];
    const i16 *filt = &rom_resamp_filt[fs_idx * 240];
    i16 *y = out + 24;
    int ii = 1 - mem_in_len, fr = 0;
    i32 m0 = st.r12k8_mem_50[0], m1 = st.r12k8_mem_50[1];
    const i32 n0 = rom_hp50_num[0], n1 = rom_hp50_num[1], n2 = rom_hp50_num[2], d0 = rom_hp50_den[0], d1 = rom_hp50_den[1];
    for (int n = 0; n < len; n++) {
        const i16 *co = &filt[fr * delay2];
        i32 acc = l_mult0(xs(ii), co[0]);
        for (int m = 1; m < delay2; m++) {
            i32 c = co[m];
            if (c) acc = l_mac0(acc, xs(ii + m), c);
        }
        i32 v = l_round(acc);
        ii += off_int;
        fr += off_frac;
        if (upfac - fr <= 0) { ii++; fr -= upfac; }
        // 2nd-order high-pass, 32-bit state (fx/resamp12k8_fx.c:88-104)
        i32 o = l_mac0(m0, n0, v);
        i32 t = l_mac0(mpy32x16k(o, d0), n1, v);
        m0 = l_add(m1, l_shl_p(t, 1));
        m1 = l_mac0(mpy32x16k(o, d1), n2, v);
        y[n] = (i16)l_round(o);
    }
    st.r12k8_mem_50[0] = m0;
    st.r12k8_mem_50[1] = m1;
    for (int k = 0; k < 24; k++) out[k] = st.r12k8_mem_out[k];
    for (int k = 0; k < 24; k++) st.r12k8_mem_out[k] = out[len + k];
}

// <a,b>/sqrt(<a,a><b,b>) in Q15 (fx/olpa_fx.c:126-155, fx/ltpf_coder_fx.c:108-141).
// sum1 and sum2 are energies plus one (>= 1): their norm is a leading-zero count, the normalising shifts cannot saturate, and
// the fractional products have a positive factor, so the -1 x -1 case of Mpy_32_32 cannot occur; the exponents are sums of a
// few values below 64.  tests/hostsim/primcheck.cpp compares this with the operator-by-operator form.
LC3_HD i32 normcorr_q15(i32 sum0, i32 sum1, i32 sum2, int sum1_prenorm, i32 sc1_in)
{
    i32 sc1 = sc1_in;
    if (!sum1_prenorm) { sc1 = clz32((u32)sum1) - 1; sum1 = (i32)((u32)sum1 << sc1); }
    const i32 sc2 = clz32((u32)sum2) - 1;
    sum2 = (i32)((u32)sum2 << sc2);
    i32 prod = (i32)(((i64)sum1 * sum2) >> 31); // both factors in [2^30, 2^31): the product is in [2^29, 2^31)
    const i32 sh = clz32((u32)prod) - 1;
    prod = (i32)((u32)prod << sh);
    i32 pe = 62 - (sc1 + sc2 + sh);
    const i32 inv = m_isqrt32(prod, &pe);       // > 0
    const i32 sc0 = l_norm(sum0);
    sum0 = (i32)((u32)sum0 << sc0);
    prod = (i32)(((i64)sum0 * inv) >> 31);
    pe = (31 - sc0) + pe;
    if (prod == 0 || s_sub(l_norm(prod), pe) >= 0) return s_max(0, l_round(l_shl(prod, pe)));
    return 32767;
}

LC3_HD i32 olpa_corr_at(const i16 *s, int lag, int n)
{
    i32 s0 = l_mult0(s[0], s[-lag]);
    i32 s1 = l_mac0(1, s[-lag], s[-lag]);
    i32 s2 = l_mac0(1, s[0], s[0]);
    for (int m = 1; m < n; m++) {
        i32 a = s[m], b = s[m - lag];
        s0 = l_mac0(s0, a, b);
        s1 = l_mac0(s1, b, b);
        s2 = l_mac0(s2, a, a);
    }
    return normcorr_q15(s0, s1, s2, 0, 0);
}

// fx/olpa_fx.c:16-222.  s12: len+1 samples; returns pitch (12.8 kHz units), *normcorr Q15.
LC3_HD i32 olpa(PitchState &st, const i16 *s12, int len, i32 *normcorr, i32 s12k8_exp, int dms)
{
    int mem_len = 114, len2 = len >> 1, acflen = len2;
    if (dms == 25) { mem_len += 16; acflen += 16; }
    i16 *mem = st.olpa_buf, *m12 = st.olpa_mem_s12k8;
    i16 *s6 = mem + mem_len, *w = s6;

    i32 sum = l_mac(l_mac(l_mult(m12[0], 4053), m12[1], 7712), m12[2], 9239);
    *w++ = (i16)l_round(l_mac(l_mac(sum, s12[0], 7712), s12[1], 4053));
    sum = l_mac(l_mac(l_mult(m12[2], 4053), s12[0], 7712), s12[1], 9239);
    *w++ = (i16)l_round(l_mac(l_mac(sum, s12[2], 7712), s12[3], 4053));
    for (int n = 5; n < len; n += 2) {
        sum = l_mac(l_mac(l_mult(s12[n - 4], 4053), s12[n - 3], 7712), s12[n - 2], 9239);
        *w++ = (i16)l_round(l_mac(l_mac(sum, s12[n - 1], 7712), s12[n], 4053));
    }
    m12[0] = s12[len - 3]; m12[1] = s12[len - 2]; m12[2] = s12[len - 1];

    // common block exponent for history and new samples
    i32 sc0 = s_sub(m_headroom16(mem, mem_len, 16), 3);
    st.olpa_exp = s_sub(st.olpa_exp, sc0);
    i32 sc1 = s_sub(m_headroom16(s6, len2, 16), 3);
    i32 s6_exp = s_sub(s12k8_exp, sc1);
    i32 sc2 = s_sub(st.olpa_exp, s6_exp);
    i32 shift;
    if (sc2 > 0) {
        m_scale_sig(s6, len2, s_sub(sc1, sc2));
        shift = sc0;
    } else {
        m_scale_sig(s6, len2, sc1);
        shift = s_add(sc0, sc2);
        st.olpa_exp = s6_exp;
    }
    if (dms == 25) s6 -= 16;
    m_scale_sig(mem, mem_len, shift);

    // weighted autocorrelation maximum over lags 17..114 (a later lag wins only if strictly
    // greater), and the unweighted maximum near the previous pitch, in one sweep
    const i32 lo = s_max(17, s_sub(st.olpa_pitch, 4)), hi = s_min(114, s_add(st.olpa_pitch, 4));
    i32 mxw = 0, mx2 = 0, p = 17, p2 = lo;
    for (int n = 17; n <= 114; n++) {
        i32 acc = l_mult0(s6[0], s6[-n]);
        for (int m = 1; m < acflen; m++) acc = l_mac0(acc, s6[m], s6[m - n]);
        i32 t = mpy32x16(acc, rom_olpa_weight[n - 17]);
        if (n == 17) mxw = t;
        else { if (l_sub(t, mxw) > 0) p = n; mxw = l_max(t, mxw); }
        if (n == lo) mx2 = acc;
        else if (n > lo && n <= hi) { if (l_sub(acc, mx2) > 0) p2 = n; mx2 = l_max(acc, mx2); }
    }
    *normcorr = olpa_corr_at(s6, p, acflen);
    if (p != p2) {
        i32 nc2 = olpa_corr_at(s6, p2, acflen);
        if (s_sub(nc2, s_mult_r(*normcorr, 27853)) > 0) { p = p2; *normcorr = nc2; }
    }
    st.olpa_pitch = p;
    for (int k = 0; k < mem_len; k++) mem[k] = mem[k + len2];
    return s_shl(p, 1);
}

// fx/ltpf_coder_fx.c:17-307.  xin: len+1 samples.  param[3] = {pitch present, active, index}; returns bits.
LC3_HD i32 ltpf_coder(PitchState &st, i32 ol_pitch, i32 ltpf_enable, int mem_in_len, i32 *param,
                      const i16 *xin, int len, i32 ol_normcorr, i32 xin_exp, int dms)
{
    i16 ac[17];
    i16 *mem = st.ltpf_buf, *x = mem + mem_in_len;
    i32 ltpf_active = 0, norm_corr = 0, ltpf_pitch, bits;

    for (int k = 0; k <= len; k++) x[k] = xin[k];

    i32 sc0 = s_sub(m_headroom16(mem, mem_in_len, 16), 3);
    st.ltpf_exp = s_sub(st.ltpf_exp, sc0);
    i32 sc1 = s_sub(m_headroom16(x, len + 1, 16), 3);
    i32 x_exp = s_sub(xin_exp, sc1);
    i32 sc2 = s_sub(st.ltpf_exp, x_exp);
    if (sc2 > 0) {
        m_scale_sig(x, len + 1, s_sub(sc1, sc2));
        m_scale_sig(mem, mem_in_len, sc0);
    } else {
        m_scale_sig(x, len + 1, sc1);
        m_scale_sig(mem, mem_in_len, s_add(sc0, sc2));
        st.ltpf_exp = x_exp;
    }

    if (s_sub(ol_normcorr, 19660) > 0) {
        const i32 min_p = s_max(s_sub(ol_pitch, 4), 32), max_p = s_min(s_add(ol_pitch, 4), 228);
        const i32 ac_lo = s_sub(min_p, 4), ac_hi = s_add(max_p, 4);
        int acflen = len;
        if (dms == 25) { acflen = len * 2; x -= len; }

        i32 sum1 = l_mac0(1, x[0], x[0]), sum2 = l_mac0(1, x[-ac_lo], x[-ac_lo]);
        for (int m = 1; m < acflen; m++) {
            sum1 = l_mac0(sum1, x[m], x[m]);
            sum2 = l_mac0(sum2, x[m - ac_lo], x[m - ac_lo]);
        }
        i32 n1 = l_norm(sum1);
        sum1 = l_shl(sum1, n1);

        for (int n = ac_lo; n <= ac_hi; n++) {
            i32 sum0 = l_mac0(0, x[0], x[-n]);
            for (int m = 1; m < acflen; m++) sum0 = l_mac0(sum0, x[m], x[m - n]);
            if (n > ac_lo) { // slide the lagged-energy window by one sample
                sum2 = l_msu0(sum2, x[acflen - 1 - (n - 1)], x[acflen - 1 - (n - 1)]);
                sum2 = l_mac0(sum2, x[-n], x[-n]);
            }
            ac[n - ac_lo] = (i16)normcorr_q15(sum0, sum1, sum2, 1, n1);
        }

        i32 ac_max = ac[min_p - ac_lo], pitch = min_p;
        for (int n = min_p + 1; n <= max_p; n++) {
            if (s_sub(ac[n - ac_lo], ac_max) > 0) pitch = n;
            ac_max = s_max(ac_max, ac[n - ac_lo]);
        }
        i32 p_int = pitch, p_fr = 0, p_idx = s_add(p_int, 283);

        if (pitch < 157) { // fractional refinement
            const int res = pitch < 127 ? 1 : 2;
            int fr0 = pitch < 127 ? -3 : -2;
            if (pitch == min_p) fr0 = 0;
            i32 best = LC3_I32_MIN;
            for (int fr = fr0; fr < 4; fr += res) {
                const i16 *h = &rom_ltpf_ac_interp[(fr + 3) * 9];
                i32 s = l_mult0(ac[p_int - ac_lo - 4], h[0]);
                for (int k = 1; k < 9; k++) s = l_mac0(s, ac[p_int - ac_lo - 4 + k], h[k]);
                if (l_sub(s, best) > 0) p_fr = fr;
                best = l_max(best, s);
            }
            if (p_fr < 0) { p_int = s_sub(p_int, 1); p_fr = s_add(p_fr, 4); }
            if (p_int >= 127) p_idx = s_add(s_add(s_shl(p_int, 1), p_fr >> 1), 126);
            else p_idx = s_sub(s_add(s_shl(p_int, 2), p_fr), 128);
        }
        ltpf_pitch = s_add(s_shl(p_int, 2), p_fr);

        const i16 *f0 = &rom_ltpf_inter_filt[0], *fp = &rom_ltpf_inter_filt[p_fr * 12];
        const i32 f00 = f0[0], f01 = f0[1], f02 = f0[2], fp0 = fp[0], fp1 = fp[1], fp2 = fp[2], fp3 = fp[3];
        i32 sum0 = 0;
        sum1 = 1; sum2 = 1;
        for (int n = 0; n < acflen; n++) {
            i32 s = l_mult(x[n + 1], f00);
            s = l_mac(s, x[n], f01);
            i32 c = s_mac_r(s, x[n - 1], f02);
            s = l_mult(x[n - p_int + 1], fp0);
            s = l_mac(s, x[n - p_int], fp1);
            s = l_mac(s, x[n - p_int - 1], fp2);
            i32 p = s_mac_r(s, x[n - p_int - 2], fp3);
            sum0 = n == 0 ? l_mult0(c, p) : l_mac0(sum0, c, p);
            sum1 = l_mac0(sum1, p, p);
            sum2 = l_mac0(sum2, c, c);
        }
        norm_corr = normcorr_q15(sum0, sum1, sum2, 0, 0);
        if (norm_corr < 0) norm_corr = 0;

        if (ltpf_enable == 1) {
            const i32 on = st.ltpf_on, mnc = st.ltpf_normcorr;
            if ((on == 0 && s_sub(mnc, 30802) > 0 && s_sub(norm_corr, 30802) > 0 &&
                 (dms == 100 || s_sub(st.ltpf_mem_normcorr, 30802) > 0)) ||
                (on == 1 && s_sub(norm_corr, 29491) > 0) ||
                (on == 1 && s_sub(s_abs(s_sub(ltpf_pitch, st.ltpf_pitch)), 8) < 0 &&
                 s_add(s_sub(norm_corr, mnc), 3277) > 0 && s_sub(norm_corr, 27525) > 0))
                ltpf_active = 1;
        }
        param[0] = 1; param[1] = ltpf_active; param[2] = p_idx;
        bits = 11;
    } else {
        norm_corr = ol_normcorr;
        param[0] = param[1] = param[2] = 0;
        bits = 1;
        ltpf_pitch = 0;
    }

    for (int n = 0; n < mem_in_len; n++) mem[n] = mem[n + len];
    st.ltpf_mem_normcorr = st.ltpf_normcorr;
    st.ltpf_normcorr = norm_corr;
    st.ltpf_on = ltpf_active;
    st.ltpf_pitch = ltpf_pitch;
    return bits;
}

// fx/attack_detector_fx.c:13-119.  10 ms frames at 32/48 kHz only (4 blocks of 40 @ 16 kHz).
// XS(i): scaled input sample i in [0, N).
template <class XS> LC3_HD void attack_detector(AttackState &st, XS xs, int N, int fs, i32 in_scaling)
{
    const int n16 = 160;
    Range rg; range_init(rg);
#pragma unroll 16
    for (int i = 0; i < N; i++) range_add(rg, xs(i));
    i32 s0 = s_add(range_headroom16(rg, 0), in_scaling);
    Range r1; range_init(r1); range_add(r1, st.filter_mem[0]); range_add(r1, st.filter_mem[1]);
    i32 s1 = s_add(range_headroom16(r1, 16), st.scaling);
    Range r2; range_init(r2); range_add(r2, st.acc_energy);
    i32 s2 = s_shr(s_add(s_add(range_headroom32(r2, 32), s_shl(st.scaling, 1)), 1), 1);
    i32 ns = s_sub(s_min(s0, s_min(s1, s2)), 2);
    i32 rs = s_sub(ns, st.scaling);
    if (rs) {
        st.filter_mem[0] = s_shl(st.filter_mem[0], rs);
        st.filter_mem[1] = s_shl(st.filter_mem[1], rs);
        st.acc_energy = l_shl(st.acc_energy, s_shl(rs, 1));
    }
    st.scaling = ns;

    const int dec = fs == 32000 ? 2 : 3;
    const i32 ds = s_sub(dec - 1, s_sub(ns, in_scaling));
    // decimate, 3-tap high-pass (on the unfiltered neighbours), block energies, in one sweep
    i32 dm2 = st.filter_mem[0], dm1 = st.filter_mem[1];
    i32 acc = st.acc_energy, detected = st.position >= 2, pos = -1, be = 0;
#pragma unroll 8
    for (int i = 0; i < n16; i++) {
        i32 d;
        if (dec == 2) d = s_add(s_shr(xs(2 * i), ds), s_shr(xs(2 * i + 1), ds));
        else d = s_add(s_shr(xs(3 * i), ds), s_add(s_shr(xs(3 * i + 1), ds), s_shr(xs(3 * i + 2), ds)));
        i32 t = l_mult(d, 12288);
        t = l_msu(t, dm1, 16384);
        t = l_mac(t, dm2, 4096);
        dm2 = dm1; dm1 = d;
        i32 f = hi16(t);
        be = l_mac(be, f, f);
        if ((i % 40) == 39) {
            i32 l = lo16(be >> 1) & 0x7fff, h = hi16(be);      // be / 8.5
            i32 q = l_mult0(l, 30840) >> 15;
            q = l_mac0(q, h, 30840) >> 2;
            if (q > acc) { pos = i / 40; detected = 1; }
            acc = l_max(acc >> 2, be);
            be = 0;
        }
    }
    st.filter_mem[0] = dm2;
    st.filter_mem[1] = dm1;
    st.acc_energy = acc;
    st.detected = detected;
    st.position = pos;
}

} // namespace lc3k
