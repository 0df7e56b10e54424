/* ORACLE (test infrastructure, not product code).
 * Spectral noise shaping: scale-factor computation (fx/sns_compute_scf_fx.c:15-164), two-stage
 * vector quantiser (fx/sns_quantize_scf_fx.c:16-550, fx/pvq_enc_fx.c:15-367, fx/pvq_dec_fx.c:55-117,
 * fx/dct2_fx.c:15-206, fx/pvq_index_fx.c:18-145,:347-507), interpolation
 * (fx/sns_interpolate_scf_fx.c:14-91) and spectrum shaping (fx/mdct_shaping_fx.c:14-35). */
#include "lc3o_internal.h"

/* fx/sns_compute_scf_fx.c:15-164.  en[] (64 entries) is expanded in place when nbands < 64. */
void o_sns_compute(i32 *en, i16 en_exp, int fs_idx, int nbands, i16 *scf, int attack, i16 att_damping)
{
    i32 d3[64];
    i16 d3e[64], d4[64];
    const i16 *pe = &rom_sns_preemph[fs_idx * 64], *pee = &rom_sns_preemph_e[fs_idx * 64];
    int i;

    if (nbands < 32) { /* each of the first (32-n) bands x4, the rest x2 */
        int k = 32 - nbands;
        for (i = nbands - 1; i >= k; i--) en[(i + k) * 2 + 1] = en[(i + k) * 2] = en[i];
        for (i = k - 1; i >= 0; i--) en[i * 4 + 3] = en[i * 4 + 2] = en[i * 4 + 1] = en[i * 4] = en[i];
    } else if (nbands < 64) { /* the first (64-n) bands x2 */
        int k = 64 - nbands;
        for (i = nbands - 1; i >= k; i--) en[i + k] = en[i];
        for (i = k - 1; i >= 0; i--) en[i * 2 + 1] = en[i * 2] = en[i];
    }

    /* 3-tap smoothing + pre-emphasis */
    i32 t = l_add(mpy32x16(en[0], 24576), l_shr(en[1], 2));
    d3[0] = mpy32x16(t, pe[0]);
    for (i = 1; i < 63; i++) {
        t = l_add(l_shr(en[i], 1), l_add(l_shr(en[i - 1], 2), l_shr(en[i + 1], 2)));
        d3[i] = mpy32x16(t, pe[i]);
    }
    t = l_add(mpy32x16(en[63], 24576), l_shr(en[62], 2));
    d3[63] = mpy32x16(t, pe[63]);
    for (i = 0; i < 64; i++) d3e[i] = s_add(en_exp, pee[i]);

    /* noise floor 40 dB below the mean */
    i16 s = d3e[63], s2 = s_add(s, 6);
    i32 mean = l_shr(d3[0], s_sub(s2, d3e[0]));
    for (i = 1; i < 64; i++) mean = l_add(mean, l_shr(d3[i], s_sub(s2, d3e[i])));
    i16 nf = o_log2_16(mean, s);
    nf = s_sub(s_max(nf, -25965), 6803);
    for (i = 0; i < 64; i++) d4[i] = s_max(nf, o_log2_16(d3[i], d3e[i]));

    /* 64 -> 16 with a 6-tap window */
    static const i16 wdn[6] = {2731, 5461, 8192, 8192, 5461, 2731};
    t = l_mult(d4[0], 8192);
    t = l_mac(t, d4[1], 8192);
    t = l_mac(t, d4[2], 8192);
    t = l_mac(t, d4[3], 5461);
    d3[0] = l_mac(t, d4[4], 2731);
    for (i = 1; i < 15; i++) {
        t = l_mult(d4[i * 4 - 1], wdn[0]);
        for (int k = 1; k < 6; k++) t = l_mac(t, d4[i * 4 - 1 + k], wdn[k]);
        d3[i] = t;
    }
    t = l_mult(d4[59], 2731);
    t = l_mac(t, d4[60], 5461);
    t = l_mac(t, d4[61], 8192);
    t = l_mac(t, d4[62], 8192);
    d3[15] = l_mac(t, d4[63], 8192);

    mean = l_shr(d3[0], 4);
    for (i = 1; i < 16; i++) mean = l_add(mean, l_shr(d3[i], 4));
    for (i = 0; i < 16; i++) scf[i] = s_mult_r(27853, l_round(l_shl(l_sub(d3[i], mean), 1)));

    if (attack) { /* 5-tap smoothing + damping.  The reference passes 32-bit sums through 16-bit
                     operator arguments (sns_compute_scf_fx.c:136-156); the truncations are kept. */
        i16 sm[16];
        i32 acc;
        sm[0] = (i16)l_shr(l_mult0((i16)l_add(l_add(scf[0], scf[1]), scf[2]), 10923), 15);
        acc = sm[0];
        sm[1] = (i16)l_shr(l_add(l_add(l_add(scf[0], scf[1]), scf[2]), scf[3]), 2);
        acc = l_add(acc, sm[1]);
        for (i = 2; i < 14; i++) {
            t = l_add(l_add(l_add(l_add(scf[i - 2], scf[i - 1]), scf[i]), scf[i + 1]), scf[i + 2]);
            sm[i] = (i16)l_shr(l_mult0((i16)t, 13107), 16);
            acc = l_add(acc, sm[i]);
        }
        sm[14] = (i16)l_shr(l_add(l_add(l_add(scf[12], scf[13]), scf[14]), scf[15]), 2);
        acc = l_add(acc, sm[14]);
        sm[15] = (i16)l_shr(l_mult0((i16)l_add(l_add(scf[13], scf[14]), scf[15]), 10923), 15);
        acc = l_add(acc, sm[15]);
        acc = l_shr(acc, 4);
        for (i = 0; i < 16; i++) scf[i] = s_mult_r(att_damping, s_sub(sm[i], (i16)acc));
    }
}

/* ------------------------------------------------------------------ stage 1: two 32x8 codebooks */
/* fx/sns_quantize_scf_fx.c:16-81: plain argmin of squared error, first minimum wins */
static int cb_search(const i16 *t, const i16 *cb)
{
    int best = 0;
    i32 bmse = 0;
    for (int r = 0; r < 32; r++) {
        i32 mse = 0;
        for (int c = 0; c < 8; c++) {
            i16 e = s_sub(cb[r * 8 + c], t[c]);
            mse = l_mac0(mse, e, e);
        }
        if (r == 0 || mse < bmse) { bmse = mse; best = r; }
    }
    return best;
}

/* ------------------------------------------------------------------ 16-point DCT-II / inverse */
static inline i16 rot(i16 x, i16 cx, i16 y, i16 cy) { return s_add(s_mult_r(x, cx), s_mult_r(y, cy)); }

/* fx/dct2_fx.c:15-109 */
static void dct16(const i16 *in, i16 *out)
{
    i16 a[16], b[16];
    int i;
    for (i = 0; i < 8; i++) a[i] = s_add(in[15 - i], in[i]);
    for (i = 10; i < 14; i++) a[i] = s_sub(in[15 - i], in[i]);
    for (i = 0; i < 4; i++) { b[i] = s_add(a[7 - i], a[i]); b[4 + i] = s_sub(a[3 - i], a[4 + i]); }
    b[8] = s_sub(in[7], in[8]); b[9] = s_sub(in[6], in[9]);
    b[10] = rot(a[10], -23170, a[13], 23170); b[11] = rot(a[11], -23170, a[12], 23170);
    b[12] = rot(a[12], 23170, a[11], 23170);  b[13] = rot(a[13], 23170, a[10], 23170);
    b[14] = s_sub(in[1], in[14]); b[15] = s_sub(in[0], in[15]);

    a[0] = s_add(b[3], b[0]); a[1] = s_add(b[2], b[1]); a[2] = s_sub(b[1], b[2]); a[3] = s_sub(b[0], b[3]);
    a[4] = b[4];
    a[5] = rot(b[5], -23170, b[6], 23170); a[6] = rot(b[6], 23170, b[5], 23170);
    a[7] = b[7];
    a[8] = s_add(b[11], b[8]); a[9] = s_add(b[10], b[9]); a[10] = s_sub(b[9], b[10]); a[11] = s_sub(b[8], b[11]);
    a[12] = s_sub(b[15], b[12]); a[13] = s_sub(b[14], b[13]); a[14] = s_add(b[13], b[14]); a[15] = s_add(b[12], b[15]);

    out[0] = rot(a[0], 8192, a[1], 8192);   out[8] = rot(a[1], -8192, a[0], 8192);
    out[4] = rot(a[2], 4433, a[3], 10703);  out[12] = rot(a[3], 4433, a[2], -10703);
    b[4] = s_add(a[5], a[4]); b[5] = s_sub(a[4], a[5]); b[6] = s_sub(a[7], a[6]); b[7] = s_add(a[6], a[7]);
    b[8] = a[8];
    b[9] = rot(a[9], -30274, a[14], 12540);  b[10] = rot(a[10], -12540, a[13], -30274);
    b[11] = a[11]; b[12] = a[12];
    b[13] = rot(a[13], 12540, a[10], -30274); b[14] = rot(a[14], 30274, a[9], 12540);
    b[15] = a[15];

    out[2] = rot(b[4], 2260, b[7], 11363);  out[10] = rot(b[5], 9633, b[6], 6436);
    out[6] = rot(b[6], 9633, b[5], -6436);  out[14] = rot(b[7], 2260, b[4], -11363);
    a[8] = s_add(b[9], b[8]); a[9] = s_sub(b[8], b[9]); a[10] = s_sub(b[11], b[10]); a[11] = s_add(b[10], b[11]);
    a[12] = s_add(b[13], b[12]); a[13] = s_sub(b[12], b[13]); a[14] = s_sub(b[15], b[14]); a[15] = s_add(b[14], b[15]);

    out[1] = rot(a[8], 1136, a[15], 11529);   out[9] = rot(a[9], 8956, a[14], 7350);
    out[5] = rot(a[10], 5461, a[13], 10217);  out[13] = rot(a[11], 11086, a[12], 3363);
    out[3] = rot(a[12], 11086, a[11], -3363); out[11] = rot(a[13], 5461, a[10], -10217);
    out[7] = rot(a[14], 8956, a[9], -7350);   out[15] = rot(a[15], 1136, a[8], -11529);
}

/* fx/dct2_fx.c:111-206 */
void o_idct16(const i16 *in, i16 *out)
{
    i16 a[16], b[16];
    a[8] = rot(in[1], 1136, in[15], -11529);  a[9] = rot(in[9], 8956, in[7], -7350);
    a[10] = rot(in[5], 5461, in[11], -10217); a[11] = rot(in[13], 11086, in[3], -3363);
    a[12] = rot(in[3], 11086, in[13], 3363);  a[13] = rot(in[11], 5461, in[5], 10217);
    a[14] = rot(in[7], 8956, in[9], 7350);    a[15] = rot(in[15], 1136, in[1], 11529);

    b[4] = rot(in[2], 2260, in[14], -11363); b[5] = rot(in[10], 9633, in[6], -6436);
    b[6] = rot(in[6], 9633, in[10], 6436);   b[7] = rot(in[14], 2260, in[2], 11363);
    b[8] = s_add(a[9], a[8]); b[9] = s_sub(a[8], a[9]); b[10] = s_sub(a[11], a[10]); b[11] = s_add(a[10], a[11]);
    b[12] = s_add(a[13], a[12]); b[13] = s_sub(a[12], a[13]); b[14] = s_sub(a[15], a[14]); b[15] = s_add(a[14], a[15]);

    a[0] = rot(in[0], 8192, in[8], 8192);   a[1] = rot(in[8], -8192, in[0], 8192);
    a[2] = rot(in[4], 4433, in[12], -10703); a[3] = rot(in[12], 4433, in[4], 10703);
    a[4] = s_add(b[5], b[4]); a[5] = s_sub(b[4], b[5]); a[6] = s_sub(b[7], b[6]); a[7] = s_add(b[6], b[7]);
    a[8] = b[8];
    a[9] = rot(b[9], -30274, b[14], 12540);  a[10] = rot(b[10], -12540, b[13], -30274);
    a[11] = b[11]; a[12] = b[12];
    a[13] = rot(b[13], 12540, b[10], -30274); a[14] = rot(b[14], 30274, b[9], 12540);
    a[15] = b[15];

    b[0] = s_add(a[3], a[0]); b[1] = s_add(a[2], a[1]); b[2] = s_sub(a[1], a[2]); b[3] = s_sub(a[0], a[3]);
    b[4] = a[4];
    b[5] = rot(a[5], -23170, a[6], 23170); b[6] = rot(a[6], 23170, a[5], 23170);
    b[7] = a[7];
    b[8] = s_add(a[11], a[8]); b[9] = s_add(a[10], a[9]); b[10] = s_sub(a[9], a[10]); b[11] = s_sub(a[8], a[11]);
    b[12] = s_sub(a[15], a[12]); b[13] = s_sub(a[14], a[13]); b[14] = s_add(a[13], a[14]); b[15] = s_add(a[12], a[15]);

    for (int i = 0; i < 4; i++) { a[i] = s_add(b[7 - i], b[i]); a[4 + i] = s_sub(b[3 - i], b[4 + i]); }
    a[10] = rot(b[10], -23170, b[13], 23170); a[11] = rot(b[11], -23170, b[12], 23170);
    a[12] = rot(b[12], 23170, b[11], 23170);  a[13] = rot(b[13], 23170, b[10], 23170);

    const i16 hi[8] = {b[15], b[14], a[13], a[12], a[11], a[10], b[9], b[8]};
    for (int i = 0; i < 8; i++) { out[i] = s_add(hi[i], a[i]); out[15 - i] = s_sub(a[i], hi[i]); }
}

/* ------------------------------------------------------------------ PVQ shape search */
typedef struct { i32 xy, yy; i16 tot; } pvq_acc;

/* fx/pvq_enc_fx.c:15-58 : floor-project |x| onto the pyramid of `num` pulses */
static void pyr_project(int dim, const i16 *xabs, i32 xsum, i16 num, i16 *y, pvq_acc *a)
{
    a->tot = 0; a->xy = 0; a->yy = 0;
    i16 sd = l_norm(xsum);
    i16 den = hi16(l_shl(xsum, sd));
    i32 ln = num;
    i16 sn = s_sub(l_norm(ln), 1);
    ln = l_shl(ln, sn);
    i16 fac = l_div(ln, den);
    i16 dsh = s_sub(sn, sd);
    for (int i = 0; i < dim; i++) {
        y[i] = hi16(l_shr(l_mult(fac, xabs[i]), dsh));
        a->tot = s_add(a->tot, y[i]);
        a->yy = l_mac0(a->yy, y[i], y[i]);
        a->xy = l_mac(a->xy, xabs[i], y[i]);
    }
}

/* fx/pvq_enc_fx.c:61-123 : add one unit pulse where corr^2/energy is largest (first max wins) */
static void add_pulse(int lo, int hi, const i16 *xabs, i16 *y, pvq_acc *a, i16 max_xabs)
{
    i16 up = l_norm(l_mac(a->xy, 1, max_xabs));
    i16 best_en = 0, best_c2 = -1;
    int imax = -1;
    for (int i = lo; i < hi; i++) {
        i16 c = l_round(l_shl(l_mac(a->xy, 1, xabs[i]), up));
        i16 c2 = s_mult(c, c);
        i16 en = lo16(l_mac(a->yy, 1, y[i]));
        if (l_msu(l_mult(c2, best_en), best_c2, en) > 0) { best_c2 = c2; best_en = en; imax = i; }
    }
    a->xy = l_mac(a->xy, xabs[imax], 1);
    a->yy = l_mac(a->yy, 1, y[imax]);
    y[imax] = s_add(y[imax], 1);
    a->tot = s_add(a->tot, 1);
}

static void put_flat(i16 *y, int dim, i16 k)
{ /* degenerate all-zero target (fx/pvq_enc_fx.c:207-226) */
    y[0] = s_shr(k, 1);
    y[dim - 1] = s_add(y[dim - 1], s_sub(k, y[0]));
}

/* fx/pvq_enc_fx.c:126-367 with the fixed 38-bit setup: 6 pulses (far), 8 (near), 10 on dims 0..9 (A),
 * 1 on dims 10..15 (B); only the far shape is seeded by projection. */
static void pvq_search(const i16 *x, i16 *y_far, i16 *y_near, i16 *yA, i16 *yB)
{
    i16 xabs[16], maxA = -1, maxB = -1;
    i32 xsum = 0, xsumA;
    int i;
    for (i = 0; i < 10; i++) { xabs[i] = s_abs(x[i]); maxA = s_max(maxA, xabs[i]); xsum = l_mac0(xsum, 1, xabs[i]); }
    xsumA = xsum;
    for (; i < 16; i++) { xabs[i] = s_abs(x[i]); maxB = s_max(maxB, xabs[i]); xsum = l_mac0(xsum, 1, xabs[i]); }
    i16 maxAll = s_max(maxA, maxB);
    memset(y_far, 0, 16 * sizeof(i16)); memset(y_near, 0, 16 * sizeof(i16));
    memset(yA, 0, 10 * sizeof(i16)); memset(yB, 0, 6 * sizeof(i16));

    if (xsum == 0) {
        put_flat(y_far, 16, 6); put_flat(y_near, 16, 8); put_flat(yA, 10, 10); put_flat(yB, 6, 1);
    } else {
        pvq_acc a;
        pyr_project(16, xabs, xsum, 6, y_far, &a);
        while (a.tot < 6) { a.yy = l_add(a.yy, 1); add_pulse(0, 16, xabs, y_far, &a, maxAll); }
        memcpy(y_near, y_far, 16 * sizeof(i16));
        while (a.tot < 8) { a.yy = l_add(a.yy, 1); add_pulse(0, 16, xabs, y_near, &a, maxAll); }

        if (xsumA == 0) {
            put_flat(yA, 10, 10);
        } else { /* restart from the near-outlier pulses that fall in section A */
            a.tot = 0; a.xy = 0; a.yy = 0;
            memcpy(yA, y_near, 10 * sizeof(i16));
            for (i = 0; i < 10; i++) {
                a.tot = s_add(a.tot, yA[i]);
                a.xy = l_mac(a.xy, xabs[i], yA[i]);
                a.yy = l_mac(a.yy, yA[i], yA[i]);
            }
            a.yy = l_shr(a.yy, 1);
            while (a.tot < 10) { a.yy = l_add(a.yy, 1); add_pulse(0, 10, xabs, yA, &a, maxA); }
        }
        int imax = 0; /* single pulse in B: last position holding the maximum */
        for (i = 10; i < 16; i++) if (xabs[i] == maxB) imax = i - 10;
        yB[imax] = 1;
    }
    for (i = 0; i < 16; i++) if (x[i] < 0) { y_far[i] = s_neg(y_far[i]); y_near[i] = s_neg(y_near[i]); }
    for (i = 0; i < 10; i++) if (x[i] < 0) yA[i] = s_neg(yA[i]);
    for (i = 0; i < 6; i++) if (x[10 + i] < 0) yB[i] = s_neg(yB[i]);
}

/* fx/pvq_dec_fx.c:67-117 : unit-energy normalisation of an integer pulse vector, Q14 */
void o_pvq_unit_norm(i16 *xq, const i16 *y, i16 kmax)
{
    i32 yy = l_mult0(y[0], y[0]);
    for (int i = 1; i < 16; i++) yy = l_mac0(yy, y[i], y[i]);
    i16 isq;
    if (l_sub(yy, 64) < 0) {
        isq = rom_isqrt_q16[yy];
    } else {
        i16 e = 15;
        i16 t = o_isqrt16(lo16(yy), &e);
        isq = s_shl(t, s_add(e, 1));
    }
    i16 sn = s_norm(kmax), st = s_sub(13, sn);
    for (int i = 0; i < 16; i++) xq[i] = l_round(l_shl(l_mult(isq, s_shl(y[i], sn)), st));
}

/* ------------------------------------------------------------------ MPVQ enumeration
 * fx/pvq_index_fx.c:18-145,:347-507 computes the offset recursion on the fly; the indices are exact
 * integers, so the table form of the same enumeration (A(n,k) = A(n-1,k)+A(n-1,k-1)+A(n,k-1)) is used. */
static u32 mpvq_A[17][12];
static int mpvq_ready;
static void mpvq_init(void)
{
    for (int k = 0; k < 12; k++) mpvq_A[0][k] = k > 0;
    for (int n = 1; n < 17; n++) {
        mpvq_A[n][0] = 0;
        mpvq_A[n][1] = 1;
        for (int k = 2; k < 12; k++) mpvq_A[n][k] = mpvq_A[n - 1][k] + mpvq_A[n - 1][k - 1] + mpvq_A[n][k - 1];
    }
    mpvq_ready = 1;
}
static void mpvq_index(const i16 *v, int dim, u32 *index, i16 *lead_sign)
{
    if (!mpvq_ready) mpvq_init();
    u32 idx = 0, next_sign = 0x80000000u;
    int kacc = 0, n = 0;
    u32 h = mpvq_A[0][0];
    for (int pos = dim - 1; pos >= 0; pos--) {
        i16 val = v[pos];
        if (!(next_sign & 0x80000000u) && val != 0) idx = 2 * idx + next_sign;
        if (val < 0) next_sign = 1;
        if (val > 0) next_sign = 0;
        idx += h;
        kacc += val < 0 ? -val : val;
        if (pos != 0) n++;
        h = mpvq_A[n][kacc];
    }
    *index = idx;
    *lead_sign = (i16)next_sign;
}

/* fx/sns_quantize_scf_fx.c:518-550 (+ :121-330 stage-2 search, :330-436 parameter layout) */
void o_sns_quantize(const i16 *scf, i32 idx[7], i16 *scf_q)
{
    i16 tgt[16], tdct[16], y[4][16], shp[4][16];
    int i;
    int iL = cb_search(scf, rom_sns_cb_lf), iH = cb_search(scf + 8, rom_sns_cb_hf);
    idx[0] = iL; idx[1] = iH;
    for (i = 0; i < 8; i++) { scf_q[i] = rom_sns_cb_lf[iL * 8 + i]; scf_q[8 + i] = rom_sns_cb_hf[iH * 8 + i]; }
    for (i = 0; i < 16; i++) tgt[i] = s_sub(scf[i], scf_q[i]);

    dct16(tgt, tdct);
    i16 yfar[16], ynear[16], yA[10], yB[6];
    pvq_search(tdct, yfar, ynear, yA, yB);
    /* shape order = spec shape_j: 0 regular (A+B), 1 regular_lf (A), 2 outlier_near, 3 outlier_far */
    for (i = 0; i < 16; i++) {
        y[0][i] = i < 10 ? yA[i] : yB[i - 10];
        y[1][i] = i < 10 ? yA[i] : 0;
        y[2][i] = ynear[i];
        y[3][i] = yfar[i];
    }
    for (int j = 0; j < 4; j++) o_pvq_unit_norm(shp[j], y[j], rom_sns_kval[j * 2]);

    i32 best = I32_MAX_;
    int bj = 0, bg = 0;
    for (int j = 0; j < 4; j++)
        for (int g = 0; g < rom_sns_gain_sz[j]; g++) {
            i32 mse = 0;
            for (i = 0; i < 16; i++) {
                i16 e = s_sub(tdct[i], s_mult_r(rom_sns_gain[j * 8 + g], shp[j][i]));
                mse = l_mac0(mse, e, e);
            }
            if (l_sub(mse, best) < 0) { best = mse; bj = j; bg = g; }
        }
    i16 rec[16];
    o_idct16(shp[bj], rec);
    for (i = 0; i < 16; i++) scf_q[i] = s_add(scf_q[i], s_mult_r(rom_sns_gain[bj * 8 + bg], rec[i]));

    u32 ix; i16 ls;
    idx[2] = bj; idx[3] = bg;
    if ((bj >> 1) == 0) {
        mpvq_index(yA, 10, &ix, &ls);
        idx[4] = ls; idx[5] = (i32)ix;
        if ((bj & 1) == 0) {
            u32 ixb; i16 lsb;
            mpvq_index(yB, 6, &ixb, &lsb);
            idx[6] = l_add(l_shl((i32)ixb, 1), lsb);
        } else idx[6] = -2;
    } else {
        mpvq_index(bj == 2 ? ynear : yfar, 16, &ix, &ls);
        idx[4] = ls; idx[5] = (i32)ix;
        idx[6] = bj == 2 ? -1 : -2;
    }
}

/* fx/sns_interpolate_scf_fx.c:14-91 (encoder side: inv_scf = 1) */
void o_sns_interp_dir(const i16 *q, i16 *g, i16 *g_exp, int nbands, int inv_scf)
{
    i16 v[64], tmp[64], t = 0, t2 = 0;
    int i;
    v[0] = v[1] = q[0];
    for (i = 1; i < 16; i++) {
        t = s_sub(q[i], q[i - 1]);
        t2 = s_mult_r(t, 8192);
        t = s_mult_r(t, 4096);
        v[4 * i - 2] = s_add(q[i - 1], t);
        v[4 * i - 1] = s_add(v[4 * i - 2], t2);
        v[4 * i] = s_add(v[4 * i - 1], t2);
        v[4 * i + 1] = s_add(v[4 * i], t2);
    }
    v[62] = s_add(v[61], t2);
    v[63] = s_add(v[62], t2);

    if (nbands < 32) {
        memcpy(tmp, v, sizeof(tmp));
        int k = 32 - nbands;
        for (i = 0; i < k; i++)
            v[i] = s_add(s_mac_r(l_mult(tmp[4 * i], 8192), tmp[4 * i + 1], 8192),
                         s_mac_r(l_mult(tmp[4 * i + 2], 8192), tmp[4 * i + 3], 8192));
        for (i = 0; i < nbands - k; i++)
            v[k + i] = s_mac_r(l_mult(tmp[4 * k + 2 * i], 16384), tmp[4 * k + 2 * i + 1], 16384);
    } else if (nbands < 64) {
        int k = 64 - nbands;
        for (i = 0; i < k; i++) v[i] = s_mac_r(l_mult(v[2 * i], 16384), v[2 * i + 1], 16384);
        for (; i < nbands; i++) v[i] = v[k + i];
    }
    for (i = 0; i < nbands; i++) g[i] = o_invlog2_16(inv_scf ? s_neg(v[i]) : v[i], &g_exp[i]);
}
void o_sns_interp(const i16 *q, i16 *g, i16 *g_exp, int nbands) { o_sns_interp_dir(q, g, g_exp, nbands, 1); }

/* fx/mdct_shaping_fx.c:14-35 */
void o_mdct_shape(i32 *d, const i16 *g, const i16 *g_exp, const i16 *bo, int nbands)
{
    int j = 0;
    for (int b = 0; b < nbands; b++)
        for (; j < bo[b + 1]; j++) d[j] = l_shl(mpy32x16(d[j], g[b]), g_exp[b]);
}
