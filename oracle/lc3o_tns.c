/* ORACLE (test infrastructure, not product code).
 * Temporal noise shaping analysis + filtering (fx/tns_coder_fx.c:21-314) with the Levinson-Durbin
 * recursion and LPC->reflection conversion (fx/levinson_fx.c:14-163). */
#include "lc3o_internal.h"

/* 1/x style helper used three times in the reference: num * (1/den) with one Newton step, <<2.
 * (fx/levinson_fx.c:45-46,:75-76,:150-152) */
static inline i32 div_newton(i32 num, i32 den)
{
    i16 inv = s_div(16383, hi16(den));
    return l_shl(mpy32x32(num, mpy32x16(l_sub(I32_MAX_, mpy32x16(den, inv)), inv)), 2);
}

/* fx/levinson_fx.c:14-116 */
static void levinson(i32 *lpc, const i32 *ac, int N, i16 *rc, i32 *pred_err)
{
    i32 tmp[9], err = ac[0], rc32;
    i16 shift = 0;
    lpc[0] = 0x8000000;
    rc32 = ac[0] != 0 ? div_newton(l_abs(ac[1]), ac[0]) : 0;
    if (ac[1] > 0) rc32 = l_neg(rc32);
    rc[0] = l_round(rc32);
    lpc[1] = l_shr(rc32, 4);
    for (int n = 2; n <= N; n++) {
        err = mpy32x32(err, l_sub(I32_MAX_, mpy32x32(rc32, rc32)));
        i16 s = l_norm(err);
        err = l_shl(err, s);
        shift = s_add(shift, s);
        i32 sum = mpy32x32(ac[1], lpc[n - 1]);
        for (int m = 2; m < n; m++) sum = l_add(sum, mpy32x32(ac[m], lpc[n - m]));
        sum = l_add(l_shl(sum, 4), ac[n]);
        rc32 = err != 0 ? div_newton(l_abs(sum), err) : 0;
        if (sum > 0) rc32 = l_neg(rc32);
        rc32 = l_shl(rc32, shift);
        rc[n - 1] = l_round(rc32);
        for (int m = 1; m < n; m++) tmp[m] = l_add(lpc[m], mpy32x32(rc32, lpc[n - m]));
        memcpy(&lpc[1], &tmp[1], (n - 1) * sizeof(i32));
        lpc[n] = l_shr(rc32, 4);
    }
    err = mpy32x32(err, l_sub(I32_MAX_, mpy32x32(rc32, rc32)));
    *pred_err = l_shr(err, shift);
}

/* fx/levinson_fx.c:119-163 */
static void lpc_to_rc(i32 *lpc, i16 *rc, int N)
{
    i32 tmp[9];
    for (int n = N; n >= 2; n--) {
        i32 rc32 = l_shl(lpc[n], 4);
        rc[n - 1] = l_round(rc32);
        i32 t0 = l_sub(I32_MAX_, l_abs(mpy32x32(rc32, rc32)));
        for (int m = 1; m < n; m++) tmp[m] = div_newton(l_sub(lpc[m], mpy32x32(lpc[n - m], rc32)), t0);
        memcpy(&lpc[1], &tmp[1], (n - 1) * sizeof(i32));
    }
    rc[0] = l_round(l_shl(lpc[1], 4));
}

/* fx/tns_coder_fx.c:288-311 : one sample through the order-`ord` FIR lattice */
static i32 lattice(int ord, const i16 *k, i32 *state, i32 x)
{
    i32 save = x;
    for (int i = 0; i < ord - 1; i++) {
        i32 t = l_add(state[i], mpy32x16(x, k[i]));
        x = l_add(x, mpy32x16(state[i], k[i]));
        state[i] = save;
        save = t;
    }
    x = l_add(x, mpy32x16(state[ord - 1], k[ord - 1]));
    state[ord - 1] = save;
    return x;
}

/* fx/tns_coder_fx.c:21-253 */
void o_tns(i16 *bits, i16 idx[16], i32 *x, int bw, i16 order[2], i16 *nfilt, int lpc_weighting, int nsub, int dms_idx)
{
    i16 buf[LC3O_MAX_N], rc[8];
    i32 rxx[9], state[8], A[9], epsP;
    int maxOrder = dms_idx == 0 ? 8 : 4;
    int xLen = rom_bw_cutoff_bin[bw] >> dms_idx;
    int startfreq[2] = {12 >> dms_idx, 0};
    const i16 *sub_lo = &rom_tns_sub_start[(dms_idx * 5 + bw) * 6], *sub_hi = &rom_tns_sub_stop[(dms_idx * 5 + bw) * 6];

    *bits = 0;
    *nfilt = 1;
    if (bw >= 3 && dms_idx <= 1) { *nfilt = 2; startfreq[1] = xLen >> 1; }
    memset(state, 0, sizeof(state));   /* lattice memory runs on from filter 0 into filter 1 */

    for (int j = 0; j < *nfilt; j++) {
        memset(rxx, 0, (maxOrder + 1) * sizeof(i32));
        for (int sd = 0; sd < nsub; sd++) {
            int lo = sub_lo[nsub * j + sd], n = sub_hi[nsub * j + sd] - lo;
            i16 headroom = o_headroom32(x + lo, n, 0), shift;
            i32 nrg = o_norm32norm(x + lo, headroom, (i16)n, &shift);
            i16 t = s_sub(l_norm(nrg), 1);
            nrg = l_shl(nrg, t);
            shift = s_sub(shift, t);
            nrg = l_add(nrg, 0x8000) & 0x7FFF0000;
            if (nrg == 0) { /* silent subdivision: unit autocorrelation, stop accumulating */
                rxx[0] = I32_MAX_;
                memset(&rxx[1], 0, maxOrder * sizeof(i32));
                break;
            }
            t = s_sub(shift, l_norm(nrg));
            t = s_shr(s_sub(1, t), 1);
            i16 shifts = s_min(t, headroom);
            i16 facs_e = s_shl(s_sub(t, shifts), 1);
            if (dms_idx != 0) facs_e = s_add(facs_e, 1);
            t = s_sub(1, s_shl(t, 1));
            nrg = l_shl(nrg, s_sub(shift, t));
            i16 facs = s_div(0x2000, l_round(nrg));

            for (int i = 0; i < n; i++) buf[i] = l_round(l_shl(x[lo + i], shifts));
            for (int lag = 0; lag <= maxOrder; lag++) {
                i32 acc = 0;
                for (int i = 0; i < n - lag; i++) acc = l_mac0(acc, buf[i], buf[i + lag]);
                if (lag != 0) acc = mpy32x32(acc, rom_tns_acf_window[lag - 1]);
                acc = l_shl(mpy32x16(acc, facs), facs_e);
                rxx[lag] = l_add(rxx[lag], acc);
            }
        }

        levinson(A, rxx, maxOrder, rc, &epsP);
        i16 sh = l_norm(epsP);
        i32 gain;
        {
            i16 inv = s_div(16383, hi16(l_shl(epsP, sh)));
            gain = mpy32x32(rxx[0], mpy32x16(l_sub(I32_MAX_, mpy32x16(l_shl(epsP, sh), inv)), inv));
        }

        if (l_sub(gain, l_shr(0x30000000, sh)) > 0) {
            if (lpc_weighting != 0 && l_sub(gain, l_shr(0x40000000, sh)) < 0) {
                i32 alpha = l_add(0x6CCCCCCD, mpy32x32(0x13333333, l_shl(l_sub(l_shl(gain, sh), 0x30000000), 3)));
                i32 w = alpha;
                for (int i = 1; i < maxOrder; i++) { A[i] = mpy32x32(A[i], w); w = mpy32x32(w, alpha); }
                A[maxOrder] = mpy32x32(A[maxOrder], w);
                lpc_to_rc(A, rc, maxOrder);
            }
            i16 *ix = &idx[8 * j];
            for (int i = 0; i < maxOrder; i++) { /* 17-level quantiser by threshold walk */
                int q = 1;
                while (q < 17 && rc[i] > rom_tns_quant_thr[q - 1]) q++;
                ix[i] = (i16)(q - 1);
            }
            int o = maxOrder - 1;
            while (o >= 0 && ix[o] == 8) o--;
            order[j] = (i16)(o + 1);

            i32 nb = rom_tns_order_bits[lpc_weighting * 8 + order[j] - 1];
            for (int i = 0; i < order[j]; i++) nb = l_add(nb, rom_tns_coef_bits[i * 17 + ix[i]]);
            *bits = s_add(*bits, s_add(2, lo16(l_shr(l_sub(nb, 1), 11))));

            for (int i = 0; i < order[j]; i++) rc[i] = rom_tns_quant_pts[ix[i]];
            int stop = (*nfilt == 2 && j == 0) ? startfreq[1] : xLen;
            for (int i = startfreq[j]; i < stop; i++) x[i] = lattice(order[j], rc, state, x[i]);
        } else {
            *bits = s_add(*bits, 1);
            order[j] = 0;
        }
    }
}
