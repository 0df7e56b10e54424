// Rate loop, per lane: global-gain estimation (fx/estimate_global_gain_fx.c:15-210), spectral
// quantiser with its context bit counter (fx/quantize_spec_fx.c:15-553, fx/ari_codec.c:1269-1308),
// gain adjustment (fx/adjust_global_gain_fx.c:14-83), residual bits (fx/residual_coding_fx.c:14-75)
// and the noise-level estimate (fx/noise_factor_fx.c:14-213).
//
// The gain estimate is split where the reference's inter-frame feedback enters: the per-quadruple
// log-energies depend only on the frame's spectrum (computed with the spectrum, frame-parallel);
// the bisection needs targetBitsOff / mem_targetBits / mem_specBits of the previous frame and runs in
// the per-stream pass.
#pragma once
#include "lc3_math.cuh"

namespace lc3k {

// gain index -> linear gain mantissa/exponent (estimate_global_gain_fx.c:200-204, adjust_...:72-75)
LC3_HD void gain_from_index(i32 gi, i32 &gain, i32 &gain_e)
{
    i32 t = l_shl_p(l_mult0(gi, 0x797D), 7);
    gain_e = s_add(lo16(t >> 25), 1);
    gain = l_round(m_invlog2(t | (i32)0xFE000000));
}

// log-energy of each quadruple, (28/20)*(7+10log10) in Q16 (estimate_global_gain_fx.c:67-110).
// Returns max |x|.
template <class A, class E> LC3_HD i32 gain_energies(A x, i32 x_e, int lg, E en)
{
    i32 x_max = 0;
    const int lg4 = lg >> 2;
    for (int i0 = 0; i0 < lg4; i0 += 2) { // two quadruples per round: eight loads in flight
        i32 w[8];
#pragma unroll
        for (int k = 0; k < 8; k++) w[k] = 4 * i0 + k < lg ? x[4 * i0 + k] : 0;
#pragma unroll
        for (int h = 0; h < 2; h++) {
            const int i = i0 + h;
            if (i >= lg4) break;
            const i32 v0 = w[4 * h], v1 = w[4 * h + 1], v2 = w[4 * h + 2], v3 = w[4 * h + 3];
            i32 m = l_max(l_max(l_abs(v0), l_abs(v1)), l_max(l_abs(v2), l_abs(v3)));
            x_max = l_max(x_max, m);
            const i32 nrm = m != 0 ? l_norm(m) : 31;
            i32 s = nrm - 2;
            // the shift leaves 2 bits of headroom, the rounded values are below 2^13 and 1 + 4 * 2 * t^2 < 2^29:
            // none of the reference's saturating operators (L_shl, round_fx, L_mac) can saturate here.  A shift by
            // nrm - 2 (which is negative for the largest values) is the exact shift by nrm followed by >> 2.
            i32 e = 1, t;
            t = (((i32)((u32)v0 << nrm) >> 2) + 0x8000) >> 16; e += 2 * t * t;
            t = (((i32)((u32)v1 << nrm) >> 2) + 0x8000) >> 16; e += 2 * t * t;
            t = (((i32)((u32)v2 << nrm) >> 2) + 0x8000) >> 16; e += 2 * t * t;
            t = (((i32)((u32)v3 << nrm) >> 2) + 0x8000) >> 16; e += 2 * t * t;
            s = s_shl(s_sub(x_e, s), 1);
            if (e == 1 && s < 0) s = 0;
            t = l_add(m_log2(e), l_shl_p(s, 25) );
            en[i] = l_add(mpy32x16k(t, 0x436E) >> 6, 0x9CCCD);
        }
    }
    return x_max;
}

struct RateState { i32 tbits_off, mem_tbits, mem_specbits; };

// second half of processEstimateGlobalGain_fx (estimate_global_gain_fx.c:44-59, :112-204)
template <class E>
LC3_HD void est_gain(E en, int lg4, i32 x_max, i32 x_e, i32 nbitsSQ, i32 qgain_off, RateState &st,
                     i32 &gain, i32 &gain_e, i32 &qgain, i32 &qgain_min)
{
    if (st.mem_tbits < 0) {
        st.tbits_off = 0;
    } else {
        i32 t = l_add(st.tbits_off, l_dep_h(s_sub(st.mem_tbits, st.mem_specbits)));
        t = l_min(40 << 16, l_max(-(40 << 16), t));
        st.tbits_off = l_add(mpy32x16k(st.tbits_off, 26214), mpy32x16k(t, 6554));
    }
    st.mem_tbits = nbitsSQ;
    nbitsSQ = s_add(nbitsSQ, l_round(st.tbits_off));

    if (x_max == 0) {
        qgain_min = qgain_off;
        qgain = 0;
        st.mem_tbits = -1;
    } else {
        x_max = l_add(m_log2(x_max), l_shl_p(x_e, 25));
        x_max = l_sub(mpy32x32k(x_max, 0x436E439A), 0xFCDD6B4);
        qgain_min = lo16(l_add(x_max, (1 << 21) + (1 << 11)) >> 21);
        qgain_min = s_max(qgain_off, s_min(s_add(255, qgain_off), qgain_min));

        i32 offset = l_dep_h(s_add(255, qgain_off));
        const i32 target = l_shl_p(l_mult(0x7D71, nbitsSQ), 1);
        i32 fac = 0x1000000;
        for (int it = 0; it < 8; it++) { // bisection on the gain index
            fac = fac >> 1;
            offset = l_sub(offset, fac);
            i32 est = 0;
            int all_zero = 1;
            for (int i = lg4 - 1; i >= 0; i--) {
                i32 t = l_sub(en[i], offset);
                if (t < 0x9CCCD) {
                    if (!all_zero) est = l_add(est, 0x3C7AE);
                } else {
                    est = l_add(est, t);
                    all_zero = 0;
                    i32 d2 = l_sub(t, 0x460000);
                    if (d2 >= 0) est = l_add(est, d2);
                }
            }
            if (est > target && !all_zero) offset = l_add(offset, fac);
        }
        i32 g = hi16(offset);
        if (g < qgain_min) st.mem_tbits = -1;
        qgain = s_sub(s_max(qgain_min, g), qgain_off);
    }
    gain_from_index(s_add(qgain, qgain_off), gain, gain_e);
}

// quantise one bin: |x|/gain + 0.375 truncated, sign restored (quantize_spec_fx.c:53-74).
// Reference form:  v = L_shl(Mpy_32_16(L_abs(x), inv_gain), s);  q = mac_r(v, -4096, 1);  sign restored.
// inv_gain comes from Inv16 and is positive, so with p = (|x| * inv_gain) >> 15 in [0, 2^31):
//   v = min(p * 2^s, 2^31 - 1)              (the only way L_shl saturates here; s <= 0 is a plain right shift)
//   q = min(v + 24576, 2^31 - 1) >> 16      (L_mac adds -8192 without saturating since v >= 0; round_fx adds 0x8000
//                                            and saturates at the top)
// which is evaluated below in 64/32-bit unsigned arithmetic.  tests/hostsim/primcheck.cpp compares it with the
// operator-by-operator form on edge values and 10^8 random inputs.
LC3_HD i32 quant_bin(i32 x, i32 inv_gain, i32 s)
{
    u32 ax = x < 0 ? 0u - (u32)x : (u32)x;
    ax = ax > (u32)LC3_I32_MAX ? (u32)LC3_I32_MAX : ax;        // L_abs saturates -2^31
    const u32 p = (u32)(((uint64_t)ax * (u32)inv_gain) >> 15);  // below 2^31 * 2^16 >> 15 = 2^32: fits 32 bits
    const int sl = imin(imax(s, 0), 32), sr = imin(imax(-s, 0), 31); // one of the two is zero
    const uint64_t w = (uint64_t)p << sl;
    u32 v = (u32)w >> sr;
    if (((u32)(w >> 32) | (v >> 31)) != 0) v = (u32)LC3_I32_MAX;  // L_shl saturated
    const u32 t = (v < (u32)LC3_I32_MAX - 24576u ? v : (u32)LC3_I32_MAX - 24576u) + 24576u; // round_fx saturates at the top
    const i32 q = (i32)(t >> 16);
    return x < 0 ? -q : q;
}

// fx/ari_codec.c:1269-1308
template <class XQ> LC3_HD i32 last_nz_pair(XQ x, int n)
{
    const int lobs[4] = {4, n >> 1, (n >> 1) + (n >> 2), (n >> 1) + (n >> 2) + (n >> 3)};
    i32 last = 0;
    int i = n;
    for (int stage = 3; stage >= 0; stage--) {
        for (; i >= lobs[stage]; i -= 2)
            if (x[i - 2] != 0 || x[i - 1] != 0) last = imax(last, i);
        if (last > 0) break;
    }
    return imax(last, 2);
}

LC3_HD i32 ari_bits(const u8 *lut, const uint16_t *bits, int t, int sym) { return bits[lut[t] * 17 + sym]; }

struct QuantOut { i32 nBits, nBits2, lastnz, lsbMode; };

// fx/quantize_spec_fx.c:15-553.  The reference unrolls three copies of the tuple loop for mode -1
// (first pass), 0 (second pass, truncation tracking) and 1 (plus LSB-plane accounting); they are one
// loop here.  xq: nt quantised bins out.  cd: 3 words per tuple (context, escape levels, MSB symbol)
// written when WRITE_CD.
template <bool WRITE_CD, class A, class XQ, class CD>
LC3_HD void quantize(QuantOut &o, A x, i32 x_e, i32 gain, i32 gain_e, XQ xq, int nt, i32 target, i32 totalBits,
                     int fs_idx, CD cd, int mode, const u8 *lut, const uint16_t *bitstab)
{
    i32 ge = gain_e;
    const i32 inv = m_inv16(gain, &ge);
    const i32 s = s_sub(s_add(x_e, ge), 15);
    for (int i = 0; i < nt; i++) xq[i] = (i16)quant_bin(x[i], inv, s);

    const i32 rateFlag = s_sub(totalBits, s_add(160, s_imult(fs_idx, 160))) > 0 ? 2 << 8 : 0;
    if (mode == 0 && s_sub(totalBits, s_add(480, s_imult(fs_idx, 160))) >= 0) mode = 1;

    const int nt_half = nt >> 1;
    i32 c = 0, nbits_lsb = 0;
    const i32 target32 = l_shl_p(target, 11);
    i32 nb = l_neg(target32), nb2 = 0;
    const i32 lastnz = last_nz_pair(xq, nt);
    i32 lastnz2 = mode >= 0 ? 2 : lastnz;

    for (int k = 0, ci = 0; k < lastnz; k += 2, ci += 3) {
        i32 t = s_add(c, rateFlag);
        if (k > nt_half) t = s_add(t, 1 << 8);
        const i32 xa = xq[k], xb = xq[k + 1];
        const i32 a1 = s_abs(xa), b1 = s_abs(xb), mx = imax(a1, b1);
        if (mx == 0) {
            if (WRITE_CD) { cd[ci] = (i16)t; cd[ci + 1] = -1; cd[ci + 2] = 0; }
            nb = l_add(nb, ari_bits(lut, bitstab, t, 0));
            c = s_add((c & 0xf) << 4, 1);
            continue; // an all-zero tuple never moves the truncation point
        }
        i32 lev, am, bm;
        if (mx < 4) {
            lev = 0; am = a1; bm = b1;
        } else if (mx < 8) {
            lev = 1; am = a1 >> 1; bm = b1 >> 1;
            nb = l_add(nb, ari_bits(lut, bitstab, t, 16));
            if (mode <= 0) nb = l_add(nb, 2 << 11);
        } else {
            lev = s_sub(13, s_norm(mx));
            for (int l = 0; l < lev; l++) nb = l_add(nb, ari_bits(lut, bitstab, t + rom_ari_esc_nb[l < 3 ? l : 3], 16));
            nb = l_add(nb, l_shl_p(mode <= 0 ? lev : s_sub(lev, 1), 12));
            am = a1 >> lev; bm = b1 >> lev;
        }
        const i32 msb = s_add(am, bm << 2);
        if (WRITE_CD) { cd[ci] = (i16)t; cd[ci + 1] = (i16)lev; cd[ci + 2] = (i16)msb; }
        nb = l_add(nb, ari_bits(lut, bitstab, t + (lev ? rom_ari_esc_nb[lev < 3 ? lev : 3] : 0), msb));
        if (mode > 0 && lev > 0) { // sign bits only for values that survive dropping the LSB
            if ((a1 >> 1) != 0) nb = l_add(nb, 1 << 11);
            if ((b1 >> 1) != 0) nb = l_add(nb, 1 << 11);
            nbits_lsb = s_add(nbits_lsb, 2);
            if (a1 == 1) nbits_lsb = s_add(nbits_lsb, 1);
            if (b1 == 1) nbits_lsb = s_add(nbits_lsb, 1);
        } else {
            if (a1 != 0) nb = l_add(nb, 1 << 11);
            if (b1 != 0) nb = l_add(nb, 1 << 11);
        }
        if (mode >= 0 && nb <= 0) { lastnz2 = k + 2; nb2 = nb; }
        if (lev == 0) c = s_add((c & 0xf) << 4, s_add(s_add(a1, b1), 1));
        else if (lev == 1) c = s_add((c & 0xf) << 4, s_add(s_shl(s_add(am, bm), 1), 1));
        else c = s_add((c & 0xf) << 4, s_add(12, imin(lev, 3)));
    }

    nb = l_add(nb, target32);
    i32 tot = s_add(lo16((l_sub(nb, 1)) >> 11), 1);
    if (mode > 0) tot = s_add(tot, nbits_lsb);
    o.nBits = tot;
    if (mode >= 0) {
        nb2 = l_add(nb2, target32);
        o.nBits2 = s_add(lo16(l_sub(nb2, 1) >> 11), 1);
    } else {
        o.nBits2 = tot;
    }
    if (mode > 0) o.nBits2 = s_add(o.nBits2, nbits_lsb);
    o.lastnz = lastnz2;
    for (int i = lastnz2; i < lastnz; i++) xq[i] = 0;
    o.lsbMode = (mode > 0 && s_sub(tot, target) > 0) ? 1 : 0;
}

// fx/adjust_global_gain_fx.c:14-83; returns 1 when the spectrum must be re-quantised
LC3_HD int adjust_gain(i32 &gg, i32 gg_min, i32 gg_off, i32 &gain, i32 &gain_e, i32 target, i32 nBits, int fs_idx)
{
    const i16 *T = rom_adjust_gain;
    i32 delta;
    if (nBits < T[0 * 5 + fs_idx]) delta = s_mult_r(s_add(nBits, 48), 2048);
    else if (nBits < T[1 * 5 + fs_idx]) delta = s_mult_r(s_add(nBits, T[4 * 5 + fs_idx]), T[3 * 5 + fs_idx]);
    else if (nBits < T[2 * 5 + fs_idx]) delta = s_mult_r(nBits, 683);
    else delta = s_mult_r(T[2 * 5 + fs_idx], 683);
    const i32 delta2 = s_add(delta, 2);

    int change = 0;
    if (gg == 255 && nBits > target) change = 1;
    if ((gg < 255 && nBits > target) || (gg > 0 && nBits < s_sub(target, delta2))) {
        if (nBits < s_sub(target, delta2)) gg = s_sub(gg, 1);
        else if (gg == 254 || nBits < s_add(target, delta)) gg = s_add(gg, 1);
        else gg = s_add(gg, 2);
        gg = s_max(gg, s_sub(gg_min, gg_off));
        gain_from_index(s_add(gg, gg_off), gain, gain_e);
        change = 1;
    }
    return change;
}

// fx/residual_coding_fx.c:14-75; bits packed LSB-first into res[] words; returns their count.
// Bins are fetched eight at a time (the loads of a group are independent, so their latencies overlap).
template <class A, class XQ>
LC3_HD i32 residual_bits(i32 x_e, A x, XQ xq, i32 gain, i32 gain_e, int L, i32 target, i32 nBits, u32 *res)
{
    i32 n = 0, m = s_add(s_sub(target, nBits), 4);
    if (m > L) m = L;
    const i32 s = s_sub(s_add(15, gain_e), x_e);
    u32 word = 0;
    bool done = false;
    for (int i0 = 0; i0 < L && !done; i0 += 8) {
        i32 q[8], xv[8];
#pragma unroll
        for (int j = 0; j < 8; j++) { const int i = imin(i0 + j, L - 1); q[j] = xq[i]; xv[j] = x[i]; }
#pragma unroll
        for (int j = 0; j < 8; j++) {
            if (done || i0 + j >= L || q[j] == 0) continue;
            u32 bit = l_sub(xv[j], l_shl(l_mult(q[j], gain), s)) >= 0;
            word |= bit << (n & 31);
            n++;
            if ((n & 31) == 0) { res[(n >> 5) - 1] = word; word = 0; }
            if (n == m) done = true;
        }
    }
    if (n & 31) res[n >> 5] = word;
    return n;
}

// mean |x| over the selected bins / gain (noise_factor_fx.c:186-204)
LC3_HD i32 level_over_gain(i32 sum, i32 cnt, i32 gg, i32 gg_e, i32 x_e, i32 norm_of, int use_norm_of)
{
    i32 s1, s2;
    i32 f = m_div3216(sum, cnt, &s1);
    f = m_div1616(f, gg, &s2);
    i32 s3 = s_sub(15, s_add(x_e, s_add(s1, s_sub(s2, gg_e))));
    s2 = s_norm(use_norm_of ? norm_of : f); // the reference normalises fac_unq1 in both halves
    if (f != 0 && s_add(s3, s2) < 0) return 32767;
    return s_shr_r(f, imin(s3, 15));
}

// fx/noise_factor_fx.c:14-213
template <class A, class XQ>
LC3_HD i32 noise_factor(i32 x_e, A x, XQ xq, i32 gg, i32 gg_e, int bw, int dms, i32 target_bytes)
{
    const int width = dms == 100 ? 3 : 1, start = dms == 100 ? 24 : (dms == 50 ? 12 : 6);
    const int N = rom_bw_cutoff_bin[bw] >> (dms == 100 ? 0 : (dms == 50 ? 1 : 2));
    const int split = target_bytes <= 20 && dms == 100;
    int c = 0, k;
    i32 nz = -2 * width - 1;
    i32 sum = 0, isum = 0;
    for (k = start - width; k < start + width; k++) nz = xq[k] != 0 ? -2 * width - 1 : s_add(nz, 1);
    // first sweep: count the selected bins, sum their |x| and (for the split variant) their indices.  Eight bins per
    // group, fetched up front (independent loads); bins from N - width on see zeros ahead of them, as in the reference
    for (int k0 = start; k0 < N; k0 += 8) {
        i32 q[8], xv[8];
        if (k0 + 8 + width <= N) { // a whole group inside the band: no bounds to check (nz and isum stay far below 2^15 / 2^31)
#pragma unroll
            for (int j = 0; j < 8; j++) { q[j] = xq[k0 + j + width]; xv[j] = x[k0 + j]; }
#pragma unroll
            for (int j = 0; j < 8; j++) {
                nz = q[j] != 0 ? -2 * width - 1 : nz + 1;
                if (nz >= 0) { c++; sum = l_add(sum, l_abs(xv[j])); isum += k0 + j; }
            }
            continue;
        }
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const int kk = k0 + j;
            q[j] = kk < N - width ? (i32)xq[kk + width] : 0;
            xv[j] = x[imin(kk, N - 1)];
        }
#pragma unroll
        for (int j = 0; j < 8; j++) {
            const int kk = k0 + j;
            if (kk >= N) break;
            nz = q[j] != 0 ? -2 * width - 1 : nz + 1;
            if (nz >= 0) { c++; sum = l_add(sum, l_abs(xv[j])); isum += kk; }
        }
    }
    i32 fac = 0;
    if (c != 0) {
        if (split) { // low/high halves around the mean index, take the minimum
            i32 s2;
            i32 mean = m_div3216(isum, c, &s2);
            mean = s_shl(mean, s2 + 1);
            i32 lo_sum = 0, hi_sum = 0;
            int lo_cnt = 0;
            nz = -2 * width - 1;
            for (k = start - width; k < start + width; k++) nz = xq[k] != 0 ? -2 * width - 1 : s_add(nz, 1);
            for (k = start; k < N; k++) {
                if (k < N - width) nz = xq[k + width] != 0 ? -2 * width - 1 : s_add(nz, 1);
                else nz = s_add(nz, 1);
                if (nz >= 0) {
                    if (k <= mean) { lo_sum = l_add(lo_sum, l_abs(x[k])); lo_cnt++; }
                    else hi_sum = l_add(hi_sum, l_abs(x[k]));
                }
            }
            i32 f1 = level_over_gain(lo_sum, lo_cnt, gg, gg_e, x_e, 0, 0);
            i32 f2 = level_over_gain(hi_sum, c - lo_cnt, gg, gg_e, x_e, f1, 1);
            fac = imin(f1, f2);
        } else {
            fac = level_over_gain(sum, c, gg, gg_e, x_e, 0, 0);
        }
    }
    i32 idx = l_round(l_sub(0x80000, l_mult(fac, 16)));
    return imin(7, imax(0, idx));
}

} // namespace lc3k
