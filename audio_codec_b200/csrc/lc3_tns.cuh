// Temporal noise shaping analysis + filtering, one frame per lane (fx/tns_coder_fx.c:21-314) with the
// Levinson-Durbin recursion and LPC -> reflection conversion (fx/levinson_fx.c:14-163).
#pragma once
#include "lc3_math.cuh"

#include "lc3_dbg.cuh"

namespace lc3k {

// f(x[lo + i]) for i = 0 .. n-1 in order, with the loads issued U at a time ahead of their use: the lane-per-frame
// passes over the spectrum are bound by the number of global loads in flight, not by bandwidth or arithmetic
template <int U, class A, class F> LC3_HD void for_batched(A x, int lo, int n, F f)
{
    int i = 0;
    for (; i + U <= n; i += U) {
        i32 v[U];
#pragma unroll
        for (int k = 0; k < U; k++) v[k] = x[lo + i + k];
#pragma unroll
        for (int k = 0; k < U; k++) f(v[k]);
    }
    for (; i < n; i++) f(x[lo + i]);
}

// num * (1/den) with one Newton step, <<2 (fx/levinson_fx.c:45-46, :75-76, :150-152)
LC3_HD i32 div_newton(i32 num, i32 den)
{
    i32 inv = s_div(16383, hi16(den));
    return l_shl_p(mpy32x32(num, mpy32x16(l_sub(LC3_I32_MAX, mpy32x16(den, inv)), inv)), 2);
}

// fx/levinson_fx.c:14-116
template <int N> LC3_HD void levinson(i32 *lpc, const i32 *ac, i16 *rc, i32 *pred_err)
{
    i32 tmp[9], err = ac[0], rc32;
    i32 shift = 0;
    lpc[0] = 0x8000000;
    rc32 = ac[0] != 0 ? div_newton(l_abs(ac[1]), ac[0]) : 0;
    if (ac[1] > 0) rc32 = l_neg(rc32);
    rc[0] = (i16)l_round(rc32);
    lpc[1] = rc32 >> 4;
#pragma unroll
    for (int n = 2; n <= N; n++) {
        err = mpy32x32(err, l_sub(LC3_I32_MAX, mpy32x32(rc32, rc32)));
        i32 s = l_norm(err);
        err = l_shl(err, s);
        shift = s_add(shift, s);
        i32 sum = mpy32x32(ac[1], lpc[n - 1]);
#pragma unroll
        for (int m = 2; m < n; m++) sum = l_add(sum, mpy32x32(ac[m], lpc[n - m]));
        sum = l_add(l_shl_p(sum, 4), ac[n]);
        rc32 = err != 0 ? div_newton(l_abs(sum), err) : 0;
        if (sum > 0) rc32 = l_neg(rc32);
        rc32 = l_shl(rc32, shift);
        rc[n - 1] = (i16)l_round(rc32);
#pragma unroll
        for (int m = 1; m < n; m++) tmp[m] = l_add(lpc[m], mpy32x32(rc32, lpc[n - m]));
#pragma unroll
        for (int m = 1; m < n; m++) lpc[m] = tmp[m];
        lpc[n] = rc32 >> 4;
    }
    err = mpy32x32(err, l_sub(LC3_I32_MAX, mpy32x32(rc32, rc32)));
    *pred_err = l_shr(err, shift);
}

// fx/levinson_fx.c:119-163
template <int N> LC3_HD void lpc_to_rc(i32 *lpc, i16 *rc)
{
    i32 tmp[9];
#pragma unroll
    for (int n = N; n >= 2; n--) {
        i32 rc32 = l_shl_p(lpc[n], 4);
        rc[n - 1] = (i16)l_round(rc32);
        i32 t0 = l_sub(LC3_I32_MAX, l_abs(mpy32x32(rc32, rc32)));
#pragma unroll
        for (int m = 1; m < n; m++) tmp[m] = div_newton(l_sub(lpc[m], mpy32x32(lpc[n - m], rc32)), t0);
#pragma unroll
        for (int m = 1; m < n; m++) lpc[m] = tmp[m];
    }
    rc[0] = (i16)l_round(l_shl_p(lpc[1], 4));
}

struct TnsOut { i32 bits, nfilt; i32 order[2]; i16 idx[16]; };

// MAXORD = 8 for 10 ms frames, 4 otherwise.  x: spectrum, filtered in place.
template <int MAXORD, class A> LC3_HD void tns_coder(TnsOut &o, A x, int bw, int lpc_weighting, int nsub, int dms_idx)
{
    i16 rc[8];
    i32 rxx[9], state[8], Ac[9], epsP;
    const int xLen = rom_bw_cutoff_bin[bw] >> dms_idx;
    int startfreq[2] = {12 >> dms_idx, 0};
    const i16 *sub_lo = &rom_tns_sub_start[(dms_idx * 5 + bw) * 6], *sub_hi = &rom_tns_sub_stop[(dms_idx * 5 + bw) * 6];

    o.bits = 0;
    o.nfilt = 1;
    o.order[0] = o.order[1] = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) o.idx[i] = 0;
    if (bw >= 3 && dms_idx <= 1) { o.nfilt = 2; startfreq[1] = xLen >> 1; }
#pragma unroll
    for (int i = 0; i < 8; i++) state[i] = 0; // lattice memory runs on from filter 0 into filter 1
    i64 state_bound = 0;                      // >= |state[i]| for every i

    for (int j = 0; j < o.nfilt; j++) {
#pragma unroll
        for (int i = 0; i <= MAXORD; i++) rxx[i] = 0;
        i64 x_bound = 0;   // >= |x[i]| over the bins of this filter, known when every subdivision was scanned
        bool scanned = true;
        for (int sd = 0; sd < nsub; sd++) {
            const int lo = sub_lo[nsub * j + sd], n = sub_hi[nsub * j + sd] - lo;
            Range rg; range_init(rg);
            for_batched<8>(x, lo, n, [&](i32 v) { range_add(rg, v); });
            x_bound = x_bound > rg.mx ? x_bound : (i64)rg.mx;
            x_bound = x_bound > -(i64)rg.mn ? x_bound : -(i64)rg.mn;
            const i32 headroom = range_headroom32(rg, 0);
            // Norm32Norm (fx/basop_util.c:534-569): energy with a sequentially adapted pre-shift
            i32 sh = headroom, nrg = 0;
            for_batched<8>(x, lo, n, [&](i32 v) {
                if (nrg >= 0x40000000) { sh = s_sub(sh, 1); nrg = nrg >> 2; }
                i32 t = l_round(l_shl_ns(v, sh)); // sh <= block headroom
                nrg += t * t;                     // nrg < 2^30 here and t^2 <= 2^30: L_mac0 cannot saturate
            });
            if (nrg != 0) {
                i32 s = s_norm(n);
                nrg = l_add(nrg, l_shl(mpy32x16(0x0000fffd, s_shl(n, s)), s_sub(15, s)));
            }
            i32 shift = s_sub(1, s_shl(sh, 1));
            i32 t = s_sub(l_norm(nrg), 1);
            nrg = l_shl(nrg, t);
            shift = s_sub(shift, t);
            nrg = l_add(nrg, 0x8000) & 0x7FFF0000;
            if (nrg == 0) { // silent subdivision: unit autocorrelation, stop accumulating
                rxx[0] = LC3_I32_MAX;
#pragma unroll
                for (int i = 1; i <= MAXORD; i++) rxx[i] = 0;
                scanned = sd == nsub - 1;
                break;
            }
            t = s_sub(shift, l_norm(nrg));
            t = s_shr(s_sub(1, t), 1);
            const i32 shifts = s_min(t, headroom);
            i32 facs_e = s_shl(s_sub(t, shifts), 1);
            if (dms_idx != 0) facs_e = s_add(facs_e, 1);
            t = s_sub(1, s_shl(t, 1));
            nrg = l_shl(nrg, s_sub(shift, t));
            const i32 facs = s_div(0x2000, l_round(nrg));

            // autocorrelation lags 0..MAXORD of the block-scaled bins, sliding window in registers
            // |partial sum of v[i] v[i-l]| <= sum v^2, so if the lag-0 energy fits 32 bits no L_mac0 of any lag can
            // have saturated and plain arithmetic is exact; otherwise redo the block with the saturating operators
            i32 acc[MAXORD + 1], hist[MAXORD + 1];
            i64 e0 = 0;
#pragma unroll
            for (int l = 0; l <= MAXORD; l++) { acc[l] = 0; hist[l] = 0; }
            for_batched<8>(x, lo, n, [&](i32 xv) {
                i32 v = l_round(l_shl_ns(xv, shifts)); // shifts <= block headroom
#pragma unroll
                for (int l = MAXORD; l > 0; l--) hist[l] = hist[l - 1];
                hist[0] = v;
                e0 += (i64)v * v;
#pragma unroll
                for (int l = 1; l <= MAXORD; l++) acc[l] += hist[l] * v;
            });
            if (e0 <= LC3_I32_MAX) {
                acc[0] = (i32)e0;
            } else {
#pragma unroll
                for (int l = 0; l <= MAXORD; l++) { acc[l] = 0; hist[l] = 0; }
                for (int i = 0; i < n; i++) {
                    i32 v = l_round(l_shl_ns(x[lo + i], shifts));
#pragma unroll
                    for (int l = MAXORD; l > 0; l--) hist[l] = hist[l - 1];
                    hist[0] = v;
#pragma unroll
                    for (int l = 0; l <= MAXORD; l++) acc[l] = l_mac0(acc[l], hist[l], v);
                }
            }
#pragma unroll
            for (int lag = 0; lag <= MAXORD; lag++) {
                i32 a = acc[lag];
                if (lag != 0) a = mpy32x32(a, rom_tns_acf_window[lag - 1]);
                a = l_shl(mpy32x16(a, facs), facs_e);
                rxx[lag] = l_add(rxx[lag], a);
            }
        }

        levinson<MAXORD>(Ac, rxx, rc, &epsP);
        const i32 sh = l_norm(epsP);
        i32 gain;
        {
            i32 ep = l_shl(epsP, sh);
            i32 inv = s_div(16383, hi16(ep));
            gain = mpy32x32(rxx[0], mpy32x16(l_sub(LC3_I32_MAX, mpy32x16(ep, inv)), inv));
        }

        if (gain > (0x30000000 >> sh)) {
            if (lpc_weighting != 0 && gain < (0x40000000 >> sh)) {
                i32 alpha = l_add(0x6CCCCCCD, mpy32x32(0x13333333, l_shl_p(l_sub(l_shl(gain, sh), 0x30000000), 3)));
                i32 w = alpha;
#pragma unroll
                for (int i = 1; i < MAXORD; i++) { Ac[i] = mpy32x32(Ac[i], w); w = mpy32x32(w, alpha); }
                Ac[MAXORD] = mpy32x32(Ac[MAXORD], w);
                lpc_to_rc<MAXORD>(Ac, rc);
            }
            i16 *ix = &o.idx[8 * j];
            int ord = 0;
#pragma unroll
            for (int i = 0; i < MAXORD; i++) { // 17-level quantiser by threshold walk
                int q = 0;
#pragma unroll
                for (int k = 0; k < 16; k++) q += rc[i] > rom_tns_quant_thr[k];
                ix[i] = (i16)q;
                if (q != 8) ord = i + 1;
            }
            o.order[j] = ord;

            i32 nb = rom_tns_order_bits[lpc_weighting * 8 + ord - 1];
            for (int i = 0; i < ord; i++) nb = l_add(nb, rom_tns_coef_bits[i * 17 + ix[i]]);
            o.bits = s_add(o.bits, s_add(2, lo16((nb - 1) >> 11)));

            i32 k[MAXORD];
#pragma unroll
            for (int i = 0; i < MAXORD; i++) k[i] = i < ord ? (i32)rom_tns_quant_pts[ix[i]] : 0;
            const int stop = (o.nfilt == 2 && j == 0) ? startfreq[1] : xLen;
            // FIR lattice of order ord (fx/tns_coder_fx.c:288-311).  Stages at and beyond ord have k = 0: they add
            // nothing to v, and the only thing to keep from them is that their memory is left alone (it runs on into
            // the second filter, whose order may be higher), so every lane walks the same eight stages without
            // branching on its own order.  Bins go through four at a time so that the loads are in flight together.
            //
            // Saturation: with |input| and |memory| <= M, a stage's two outputs are bounded by M (1 + |k|) + 1 (the
            // product is rounded down), so M grows stage by stage to a bound on every value the filter can form.
            // The subdivisions scanned above tile the filter's bins (tests/test_rom_bounds.py) and give M's start.
            // The shaped spectrum has about nine bits of headroom, so the bound stays below 2^31 in practice and the
            // L_add of the reference is a plain addition; otherwise the saturating form runs.  k is never -32768
            // (tests/test_rom_bounds.py), so Mpy_32_16's one saturating case cannot occur either.
            i64 M = x_bound > state_bound ? x_bound : state_bound;
#pragma unroll
            for (int i = 0; i < MAXORD; i++) M += ((M * (k[i] < 0 ? -k[i] : k[i])) >> 15) + 2;
            const bool plain = scanned && M <= LC3_I32_MAX;
            state_bound = plain ? M : (i64)LC3_I32_MAX + 1;
            auto run = [&](auto sat) {
                constexpr bool SAT = decltype(sat)::value;
                auto lattice = [&](i32 v) -> i32 {
                    i32 save = v;
#pragma unroll
                    for (int s = 0; s < MAXORD; s++) {
                        const i32 st = state[s];
                        const i32 p = mpy32x16k(v, k[s]), q = mpy32x16k(st, k[s]);
                        const i32 t = SAT ? l_add(st, p) : add_proved(st, p);
                        v = SAT ? l_add(v, q) : add_proved(v, q);
                        state[s] = s < ord ? save : st;
                        save = t; // not used past stage ord - 1
                    }
                    return v;
                };
                int i = startfreq[j];
                for (; i + 4 <= stop; i += 4) {
                    i32 v0 = x[i], v1 = x[i + 1], v2 = x[i + 2], v3 = x[i + 3];
                    v0 = lattice(v0); v1 = lattice(v1); v2 = lattice(v2); v3 = lattice(v3);
                    x[i] = v0; x[i + 1] = v1; x[i + 2] = v2; x[i + 3] = v3;
                }
                for (; i < stop; i++) x[i] = lattice(x[i]);
            };
            if (plain) run(std::false_type()); else run(std::true_type());
        } else {
            o.bits = s_add(o.bits, 1);
            o.order[j] = 0;
        }
    }
}

} // namespace lc3k
