/* ORACLE (test infrastructure, not product code).
 * Complex FFT of the DCT-IV kernel for the lengths the encoder uses (10,20,30,40,60,80,120,160,240),
 * restating fx/fft.c: prime-factor leaves fft5/fft10/fft15/fft20/fft30 (:204-2598), the 5x8
 * routine fft40 (:3208-3362) and the generic "leaf x radix-4/8" combiner fftN2 (:3364-3651),
 * dispatched as in BASOP_cfft (:3668-3775).  The reference spells every butterfly out; here the
 * Good-Thomas index maps are written as modular arithmetic and the butterflies as loops, with the
 * same operation order, shifts and rounding.  Data is interleaved re,im.
 */
#include "lc3o_internal.h"

typedef struct { i32 re, im; } cpx;

/* 16-bit constants = WORD322WORD16 of the 32-bit ones (fx/fft.c:51-81) */
static const i32 K5_16[5] = {31164, -25212, -11904, 18318, -20480};                 /* C51..C55 */
static const i32 K5_32[5] = {0x79bc3854, (i32)0x9d839db0, (i32)0xd18053ce, 0x478dde64, (i32)0xb0000001};
#define K31 ((i16)-28378)
#define K81 ((i16)23170)
#define K81_32 0x5a82799a
#define K82_32 ((i32)0xa57d8666)

static inline i32 mulk(i32 x, i32 k, int wide) { return wide ? mpy32x32(x, k) : mpy32x16(x, (i16)k); }

/* one real-valued half of the 5-point Winograd kernel (fx/fft.c:221-243 and every "FFT5 stage") */
static void half5(i32 x0, i32 a1, i32 a2, i32 a3, i32 a4, const i32 *K, int wide, i32 *y0, i32 r[5])
{
    i32 r1 = l_add(a1, a4), r4 = l_sub(a1, a4), r3 = l_add(a2, a3), r2 = l_sub(a2, a3);
    i32 t = mulk(l_sub(r1, r3), K[3], wide);
    r1 = l_add(r1, r3);
    *y0 = l_add(x0, r1);
    r1 = l_add(*y0, l_shl(mulk(r1, K[4], wide), 1));
    r3 = l_sub(r1, t);
    r1 = l_add(r1, t);
    t = mulk(l_add(r4, r2), K[0], wide);
    r4 = l_add(t, l_shl(mulk(r4, K[1], wide), 1));
    r2 = l_add(t, mulk(r2, K[2], wide));
    r[1] = r1; r[2] = r2; r[3] = r3; r[4] = r4;
}

/* complex 5-point DFT, inputs a[0..4], outputs o[0..4] in natural order */
static void dft5(const cpx *a, cpx *o, int wide)
{
    const i32 *K = wide ? K5_32 : K5_16;
    i32 r[5], s[5];
    half5(a[0].re, a[1].re, a[2].re, a[3].re, a[4].re, K, wide, &o[0].re, r);
    half5(a[0].im, a[1].im, a[2].im, a[3].im, a[4].im, K, wide, &o[0].im, s);
    o[1].re = l_add(r[1], s[2]); o[4].re = l_sub(r[1], s[2]);
    o[2].re = l_sub(r[3], s[4]); o[3].re = l_add(r[3], s[4]);
    o[1].im = l_sub(s[1], r[2]); o[4].im = l_add(s[1], r[2]);
    o[2].im = l_add(s[3], r[4]); o[3].im = l_sub(s[3], r[4]);
}

/* complex 3-point DFT (fx/fft.c "FFT3 stage"): o0, o_minus, o_plus */
static void dft3(cpx a0, cpx a1, cpx a2, cpx *o0, cpx *om, cpx *op)
{
    i32 r1 = l_add(a1.re, a2.re), r2 = mpy32x16(l_sub(a1.re, a2.re), K31);
    o0->re = l_add(a0.re, r1);
    r1 = l_sub(a0.re, l_shr(r1, 1));
    i32 s1 = l_add(a1.im, a2.im), s2 = mpy32x16(l_sub(a1.im, a2.im), K31);
    o0->im = l_add(a0.im, s1);
    s1 = l_sub(a0.im, l_shr(s1, 1));
    om->re = l_sub(r1, s2); op->re = l_add(r1, s2);
    om->im = l_add(s1, r2); op->im = l_sub(s1, r2);
}

/* complex 4-point DFT (fx/fft.c:153-201 and the "FFT4 stage"s) */
static void dft4(const cpx *c, cpx *o)
{
    i32 t0 = l_add(c[0].re, c[2].re), t2 = l_sub(c[0].re, c[2].re);
    i32 t1 = l_add(c[0].im, c[2].im), t3 = l_sub(c[0].im, c[2].im);
    i32 t4 = l_add(c[1].re, c[3].re), t7 = l_sub(c[1].re, c[3].re);
    i32 t5 = l_add(c[3].im, c[1].im), t6 = l_sub(c[3].im, c[1].im);
    o[0].re = l_add(t0, t4); o[0].im = l_add(t1, t5);
    o[1].re = l_sub(t2, t6); o[1].im = l_sub(t3, t7);
    o[2].re = l_sub(t0, t4); o[2].im = l_sub(t1, t5);
    o[3].re = l_add(t2, t6); o[3].im = l_add(t3, t7);
}

static inline cpx ldsh(const i32 *x, int i, int sh)
{
    cpx c = {l_shr(x[2 * i], (i16)sh), l_shr(x[2 * i + 1], (i16)sh)};
    return c;
}
static inline void st(i32 *x, int i, cpx c) { x[2 * i] = c.re; x[2 * i + 1] = c.im; }

/* 15 = 3 x 5 prime-factor core on already-scaled inputs a[] -> z[] natural order (fx/fft.c:611-915) */
static void core15(const cpx *a, cpx *z)
{
    cpx y[3][5], in[5];
    for (int k = 0; k < 3; k++) {
        for (int j = 0; j < 5; j++) in[j] = a[(5 * k + 3 * j) % 15];
        dft5(in, y[k], 0);
    }
    for (int j = 0; j < 5; j++)
        dft3(y[0][j], y[1][j], y[2][j], &z[(6 * j) % 15], &z[(6 * j + 10) % 15], &z[(6 * j + 5) % 15]);
}

static void fft10(i32 *x)
{ /* fx/fft.c:433-596 : 2 x 5 */
    cpx in[5], y[2][5];
    for (int k = 0; k < 2; k++) {
        for (int j = 0; j < 5; j++) in[j] = ldsh(x, (5 * k + 6 * j) % 10, 5);
        dft5(in, y[k], 0);
    }
    for (int j = 0; j < 5; j++) {
        cpx p = {l_add(y[0][j].re, y[1][j].re), l_add(y[0][j].im, y[1][j].im)};
        cpx m = {l_sub(y[0][j].re, y[1][j].re), l_sub(y[0][j].im, y[1][j].im)};
        st(x, 2 * j, p);
        st(x, (2 * j + 5) % 10, m);
    }
}

static void fft15(i32 *x)
{ /* fx/fft.c:611-915 */
    cpx a[15], z[15];
    for (int i = 0; i < 15; i++) a[i] = ldsh(x, i, 5);
    core15(a, z);
    for (int i = 0; i < 15; i++) st(x, i, z[i]);
}

static void fft20(i32 *x)
{ /* fx/fft.c:1395-1775 : 4 x 5 */
    cpx in[5], y[4][5], c[4], o[4];
    for (int k = 0; k < 4; k++) {
        for (int j = 0; j < 5; j++) in[j] = ldsh(x, (5 * k + 16 * j) % 20, 5);
        dft5(in, y[k], 0);
    }
    for (int j = 0; j < 5; j++) {
        for (int k = 0; k < 4; k++) c[k] = y[k][j];
        dft4(c, o);
        for (int m = 0; m < 4; m++) st(x, (4 * j + 5 * m) % 20, o[m]);
    }
}

static void fft30(i32 *x)
{ /* fx/fft.c:1791-2598 : 2 x 15 */
    cpx a[15], z[2][15];
    for (int h = 0; h < 2; h++) {
        for (int i = 0; i < 15; i++) a[i] = ldsh(x, (16 * i + 15 * h) % 30, 5);
        core15(a, z[h]);
    }
    for (int i = 0; i < 15; i++) {
        int c = (8 * i) % 15;
        i32 r1 = l_shr(z[0][c].re, 1), r2 = l_shr(z[1][c].re, 1);
        i32 r3 = l_shr(z[0][c].im, 1), r4 = l_shr(z[1][c].im, 1);
        cpx p = {l_add(r1, r2), l_add(r3, r4)}, m = {l_sub(r1, r2), l_sub(r3, r4)};
        if (i & 1) { st(x, 15 + i, p); st(x, i, m); }
        else       { st(x, i, p); st(x, 15 + i, m); }
    }
}

/* radix-8 column butterfly with the interleaved >>2 scaling (fx/fft.c:3293-3357, :3499-3587) */
static void dft8(const cpx *v, cpx *o, int wide)
{
    i32 t00 = l_shr(l_add(v[0].re, v[4].re), 2), t02 = l_shr(l_sub(v[0].re, v[4].re), 2);
    i32 t01 = l_shr(l_add(v[0].im, v[4].im), 2), t03 = l_shr(l_sub(v[0].im, v[4].im), 2);
    i32 t04 = l_shr(l_add(v[1].re, v[5].re), 2), t06 = l_sub(v[1].re, v[5].re);
    i32 t05 = l_shr(l_add(v[1].im, v[5].im), 2), t07 = l_sub(v[1].im, v[5].im);
    i32 t08 = l_shr(l_add(v[2].re, v[6].re), 2), t10 = l_shr(l_sub(v[2].re, v[6].re), 2);
    i32 t09 = l_shr(l_add(v[2].im, v[6].im), 2), t11 = l_shr(l_sub(v[2].im, v[6].im), 2);
    i32 t12 = l_shr(l_add(v[3].re, v[7].re), 2), t14 = l_sub(v[3].re, v[7].re);
    i32 t13 = l_shr(l_add(v[3].im, v[7].im), 2), t15 = l_sub(v[3].im, v[7].im);

    i32 s00 = l_add(t00, t08), s04 = l_sub(t00, t08), s01 = l_add(t01, t09), s05 = l_sub(t01, t09);
    i32 s08 = l_sub(t02, t11), s10 = l_add(t02, t11), s09 = l_add(t03, t10), s11 = l_sub(t03, t10);
    i32 s02 = l_add(t04, t12), s07 = l_sub(t04, t12), s03 = l_add(t05, t13), s06 = l_sub(t13, t05);

    i32 u1 = l_shr(l_add(t06, t14), 2), u2 = l_shr(l_sub(t06, t14), 2);
    i32 u0 = l_shr(l_add(t07, t15), 2), u3 = l_shr(l_sub(t07, t15), 2);
    i32 kp = wide ? K81_32 : K81, kn = wide ? K82_32 : -K81;
    i32 s12 = mulk(l_add(u0, u2), kp, wide), s14 = mulk(l_sub(u0, u2), kp, wide);
    i32 s13 = mulk(l_sub(u3, u1), kp, wide), s15 = mulk(l_add(u1, u3), kn, wide);

    o[0].re = l_add(s00, s02); o[0].im = l_add(s01, s03);
    o[1].re = l_add(s10, s12); o[1].im = l_add(s11, s13);
    o[2].re = l_sub(s04, s06); o[2].im = l_sub(s05, s07);
    o[3].re = l_add(s08, s14); o[3].im = l_add(s09, s15);
    o[4].re = l_sub(s00, s02); o[4].im = l_sub(s01, s03);
    o[5].re = l_sub(s10, s12); o[5].im = l_sub(s11, s13);
    o[6].re = l_add(s04, s06); o[6].im = l_add(s05, s07);
    o[7].re = l_sub(s08, s14); o[7].im = l_sub(s09, s15);
}

static void leaf(i32 *x, int n)
{
    switch (n) {
    case 10: fft10(x); break;
    case 15: fft15(x); break;
    case 20: fft20(x); break;
    case 30: fft30(x); break;
    }
}

/* fx/fft.c:3364-3651 : dim2 rows of a dim1-point leaf, twiddle, dim1 columns of radix dim2 (4|8) */
static void fft_n2(i32 *x, const i16 *W, int dim1, int dim2, int sc, int Woff)
{
    i32 w[2 * 240];
    cpx v[8], o[8];
    for (int i = 0; i < dim2; i++)
        for (int j = 0; j < dim1; j++) {
            w[2 * (i * dim1 + j)] = x[2 * (i + j * dim2)];
            w[2 * (i * dim1 + j) + 1] = x[2 * (i + j * dim2) + 1];
        }
    for (int i = 0; i < dim2; i++) leaf(&w[2 * i * dim1], dim1);

    int sh = dim2 == 4 ? 2 : 1;              /* SCALEFACTOR60-SCALEFACTOR15 | 1 */
    int tws = dim2 == 4 ? 8 : sc;            /* twiddle row stride factor (fx/fft.c:3440 j=8) */
    for (int i = 0; i < dim1; i++) {
        for (int k = 0; k < dim2; k++) {
            i32 a = w[2 * (i + k * dim1)], b = w[2 * (i + k * dim1) + 1];
            if (k == 0 || i == 0) {
                v[k].re = l_shr(a, (i16)sh);
                v[k].im = l_shr(b, (i16)sh);
            } else {
                i16 c = W[sc * i + tws * k * dim1 - Woff], d = W[sc * i + tws * k * dim1 + 1 - Woff];
                v[k].re = l_shr(l_sub(mpy32x16(a, c), mpy32x16(b, d)), (i16)sh);
                v[k].im = l_shr(l_add(mpy32x16(a, d), mpy32x16(b, c)), (i16)sh);
            }
        }
        if (dim2 == 4) dft4(v, o); else dft8(v, o, 0);
        for (int m = 0; m < dim2; m++) st(x, i + m * dim1, o[m]);
    }
}

/* fx/fft.c:3208-3362 : 8 rows of fft5 (32-bit constants), conjugate twiddles, 5 radix-8 columns */
static void fft40(i32 *x)
{
    cpx w[40], in[5], v[8], o[8];
    for (int i = 0; i < 8; i++) {
        for (int j = 0; j < 5; j++) in[j] = ldsh(x, i + j * 8, 4);
        dft5(in, &w[i * 5], 1);
    }
    for (int i = 0; i < 5; i++) {
        for (int k = 0; k < 8; k++) {
            cpx a = w[i + k * 5];
            if (k == 0 || i == 0) {
                v[k].re = l_shr(a.re, 1);
                v[k].im = l_shr(a.im, 1);
            } else {
                i32 c = rom_rot40[2 * (i - 1) + (k - 1) * 8], d = rom_rot40[2 * (i - 1) + (k - 1) * 8 + 1];
                v[k].re = l_shr(l_add(mpy32x32(a.re, c), mpy32x32(a.im, d)), 1);
                v[k].im = l_shr(l_sub(mpy32x32(a.im, c), mpy32x32(a.re, d)), 1);
            }
        }
        dft8(v, o, 1);
        for (int m = 0; m < 8; m++) st(x, i + m * 5, o[m]);
    }
}

/* fx/fft.c:3668-3775 */
void o_cfft(i32 *x, int len, i16 *scale)
{
    int sf = 0;
    switch (len) {
    case 10: fft10(x); sf = 5; break;
    case 20: fft20(x); sf = 5; break;
    case 30: fft30(x); sf = 6; break;
    case 40: fft40(x); sf = 7; break;
    case 60: fft_n2(x, rom_rot480, 15, 4, 4, 60); sf = 7; break;
    case 80: fft_n2(x, rom_rot320, 10, 8, 4, 40); sf = 8; break;
    case 120: fft_n2(x, rom_rot480, 15, 8, 4, 60); sf = 8; break;
    case 160: fft_n2(x, rom_rot320, 20, 8, 2, 40); sf = 8; break;
    case 240: fft_n2(x, rom_rot480, 30, 8, 2, 60); sf = 9; break;
    }
    *scale = s_add(*scale, (i16)sf);
}
