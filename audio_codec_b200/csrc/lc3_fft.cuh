// Complex FFT of the DCT-IV kernel, lengths 10,20,30,40,60,80,120,160,240 (fx/fft.c).
//
// One lane transforms one frame; all index maps are compile-time so the leaf transforms
// (prime-factor 5x2, 5x3, 5x4, 15x2: fx/fft.c:204-2598) live entirely in registers.  The composite
// lengths follow fftN2 (fx/fft.c:3364-3651): DIM2 leaf transforms of length LEAF read straight from
// the strided input into a scratch row, then LEAF radix-4/8 column butterflies with 16-bit twiddles
// write the result back.  fft40 (fx/fft.c:3208-3362) is the same shape with 32-bit constants and
// conjugated twiddles.  Every shift, product and saturating add is in the reference's order.
#pragma once
#include "lc3_rom.cuh"

namespace lc3k {

// ---- additions of the FFT network.
// The reference uses saturating L_add / L_sub / L_shl throughout (fx/fft.c).  The only caller is the DCT-IV of the MDCT
// (lc3_mdct.cuh, lc3_mdct_tile.cuh), whose pre-twiddle bounds every input component by 0.7072 * 2^31.  Under that bound
// none of these operations can saturate: tests/hostsim/primcheck.cpp pushes every basis input through each FFT length,
// records every intermediate (the network is linear and data independent), and checks that the sum of absolute
// contributions of all inputs at their maximum, plus rounding slack, stays below 2^31 at every operation.  That check
// closes for every length except 160 (N = 320: largest sum of |coefficients| 1.63, times 0.7072 > 1), which therefore
// keeps the saturating operators (SAT = true); all other lengths use plain (wrapping) arithmetic.  -DLC3_FFT_TRACE turns
// the three into recording hooks for the check.
// FFT lengths whose bound check does NOT close (see below): they keep the reference's saturating operators.
template <int LEN> struct FftNeedsSat { static constexpr bool v = LEN == 160; };
#if defined(LC3_FFT_TRACE) && !defined(__CUDA_ARCH__)
struct FftTrace { static inline i64 *acc = nullptr; static inline int n = 0, cap = 0; };
inline i32 fft_rec(i64 v)
{
    if (FftTrace::acc && FftTrace::n < FftTrace::cap) FftTrace::acc[FftTrace::n] += v < 0 ? -v : v;
    FftTrace::n++;
    return (i32)v;
}
template <bool SAT> inline i32 fa(i32 a, i32 b) { return fft_rec((i64)a + b); }
template <bool SAT> inline i32 fs(i32 a, i32 b) { return fft_rec((i64)a - b); }
template <bool SAT> inline i32 fsh(i32 a, i32 n) { return fft_rec((i64)a * ((i64)1 << n)); }
#else
template <bool SAT> LC3_HD i32 fa(i32 a, i32 b) { return SAT ? l_add(a, b) : (i32)((u32)a + (u32)b); }
template <bool SAT> LC3_HD i32 fs(i32 a, i32 b) { return SAT ? l_sub(a, b) : (i32)((u32)a - (u32)b); }
template <bool SAT> LC3_HD i32 fsh(i32 a, i32 n) { return SAT ? l_shl_p(a, n) : (i32)((u32)a << n); }
#endif
// column twiddle product: without the bound the -1 x -1 case of Mpy_32_16 has to be kept
template <bool SAT> LC3_HD i32 mul16(i32 x, i32 y) { return SAT ? mpy32x16(x, y) : mpy32x16k(x, y); }

struct cpx { i32 re, im; };

template <bool WIDE> LC3_HD i32 mulk(i32 x, i32 k) { return WIDE ? mpy32x32k(x, k) : mpy32x16k(x, k); }

template <bool WIDE> struct K5 {};
template <> struct K5<false> { // C51..C55 as 16-bit (fx/fft.c:51-81)
    static constexpr i32 k0 = 31164, k1 = -25212, k2 = -11904, k3 = 18318, k4 = -20480;
};
template <> struct K5<true> {
    static constexpr i32 k0 = 0x79bc3854, k1 = (i32)0x9d839db0, k2 = (i32)0xd18053ce, k3 = 0x478dde64, k4 = (i32)0xb0000001;
};

// one real-valued half of the 5-point Winograd kernel (fx/fft.c:221-243)
template <bool WIDE, bool SAT> LC3_HD void half5(i32 x0, i32 a1, i32 a2, i32 a3, i32 a4, i32 &y0, i32 &o1, i32 &o2, i32 &o3, i32 &o4)
{
    typedef K5<WIDE> K;
    i32 r1 = fa<SAT>(a1, a4), r4 = fs<SAT>(a1, a4), r3 = fa<SAT>(a2, a3), r2 = fs<SAT>(a2, a3);
    i32 t = mulk<WIDE>(fs<SAT>(r1, r3), K::k3);
    r1 = fa<SAT>(r1, r3);
    y0 = fa<SAT>(x0, r1);
    r1 = fa<SAT>(y0, fsh<SAT>(mulk<WIDE>(r1, K::k4), 1));
    r3 = fs<SAT>(r1, t);
    r1 = fa<SAT>(r1, t);
    t = mulk<WIDE>(fa<SAT>(r4, r2), K::k0);
    r4 = fa<SAT>(t, fsh<SAT>(mulk<WIDE>(r4, K::k1), 1));
    r2 = fa<SAT>(t, mulk<WIDE>(r2, K::k2));
    o1 = r1; o2 = r2; o3 = r3; o4 = r4;
}

template <bool WIDE, bool SAT> LC3_HD void dft5(const cpx *a, cpx *o)
{
    i32 r1, r2, r3, r4, s1, s2, s3, s4;
    half5<WIDE, SAT>(a[0].re, a[1].re, a[2].re, a[3].re, a[4].re, o[0].re, r1, r2, r3, r4);
    half5<WIDE, SAT>(a[0].im, a[1].im, a[2].im, a[3].im, a[4].im, o[0].im, s1, s2, s3, s4);
    o[1].re = fa<SAT>(r1, s2); o[4].re = fs<SAT>(r1, s2);
    o[2].re = fs<SAT>(r3, s4); o[3].re = fa<SAT>(r3, s4);
    o[1].im = fs<SAT>(s1, r2); o[4].im = fa<SAT>(s1, r2);
    o[2].im = fa<SAT>(s3, r4); o[3].im = fs<SAT>(s3, r4);
}

// 3-point DFT ("FFT3 stage"): outputs o0, o_minus, o_plus
template <bool SAT> LC3_HD void dft3(cpx a0, cpx a1, cpx a2, cpx &o0, cpx &om, cpx &op)
{
    const i32 K31 = -28378;
    i32 r1 = fa<SAT>(a1.re, a2.re), r2 = mpy32x16k(fs<SAT>(a1.re, a2.re), K31);
    o0.re = fa<SAT>(a0.re, r1);
    r1 = fs<SAT>(a0.re, r1 >> 1);
    i32 s1 = fa<SAT>(a1.im, a2.im), s2 = mpy32x16k(fs<SAT>(a1.im, a2.im), K31);
    o0.im = fa<SAT>(a0.im, s1);
    s1 = fs<SAT>(a0.im, s1 >> 1);
    om.re = fs<SAT>(r1, s2); op.re = fa<SAT>(r1, s2);
    om.im = fa<SAT>(s1, r2); op.im = fs<SAT>(s1, r2);
}

// 4-point DFT (fx/fft.c:153-201)
template <bool SAT> LC3_HD void dft4(const cpx *c, cpx *o)
{
    i32 t0 = fa<SAT>(c[0].re, c[2].re), t2 = fs<SAT>(c[0].re, c[2].re);
    i32 t1 = fa<SAT>(c[0].im, c[2].im), t3 = fs<SAT>(c[0].im, c[2].im);
    i32 t4 = fa<SAT>(c[1].re, c[3].re), t7 = fs<SAT>(c[1].re, c[3].re);
    i32 t5 = fa<SAT>(c[3].im, c[1].im), t6 = fs<SAT>(c[3].im, c[1].im);
    o[0].re = fa<SAT>(t0, t4); o[0].im = fa<SAT>(t1, t5);
    o[1].re = fs<SAT>(t2, t6); o[1].im = fs<SAT>(t3, t7);
    o[2].re = fs<SAT>(t0, t4); o[2].im = fs<SAT>(t1, t5);
    o[3].re = fa<SAT>(t2, t6); o[3].im = fa<SAT>(t3, t7);
}

// radix-8 column butterfly with interleaved >>2 scaling (fx/fft.c:3293-3357, :3499-3587)
template <bool WIDE, bool SAT> LC3_HD void dft8(const cpx *v, cpx *o)
{
    i32 t00 = fa<SAT>(v[0].re, v[4].re) >> 2, t02 = fs<SAT>(v[0].re, v[4].re) >> 2;
    i32 t01 = fa<SAT>(v[0].im, v[4].im) >> 2, t03 = fs<SAT>(v[0].im, v[4].im) >> 2;
    i32 t04 = fa<SAT>(v[1].re, v[5].re) >> 2, t06 = fs<SAT>(v[1].re, v[5].re);
    i32 t05 = fa<SAT>(v[1].im, v[5].im) >> 2, t07 = fs<SAT>(v[1].im, v[5].im);
    i32 t08 = fa<SAT>(v[2].re, v[6].re) >> 2, t10 = fs<SAT>(v[2].re, v[6].re) >> 2;
    i32 t09 = fa<SAT>(v[2].im, v[6].im) >> 2, t11 = fs<SAT>(v[2].im, v[6].im) >> 2;
    i32 t12 = fa<SAT>(v[3].re, v[7].re) >> 2, t14 = fs<SAT>(v[3].re, v[7].re);
    i32 t13 = fa<SAT>(v[3].im, v[7].im) >> 2, t15 = fs<SAT>(v[3].im, v[7].im);

    i32 s00 = fa<SAT>(t00, t08), s04 = fs<SAT>(t00, t08), s01 = fa<SAT>(t01, t09), s05 = fs<SAT>(t01, t09);
    i32 s08 = fs<SAT>(t02, t11), s10 = fa<SAT>(t02, t11), s09 = fa<SAT>(t03, t10), s11 = fs<SAT>(t03, t10);
    i32 s02 = fa<SAT>(t04, t12), s07 = fs<SAT>(t04, t12), s03 = fa<SAT>(t05, t13), s06 = fs<SAT>(t13, t05);

    i32 u1 = fa<SAT>(t06, t14) >> 2, u2 = fs<SAT>(t06, t14) >> 2;
    i32 u0 = fa<SAT>(t07, t15) >> 2, u3 = fs<SAT>(t07, t15) >> 2;
    const i32 kp = WIDE ? 0x5a82799a : 23170, kn = WIDE ? (i32)0xa57d8666 : -23170;
    i32 s12 = mulk<WIDE>(fa<SAT>(u0, u2), kp), s14 = mulk<WIDE>(fs<SAT>(u0, u2), kp);
    i32 s13 = mulk<WIDE>(fs<SAT>(u3, u1), kp), s15 = mulk<WIDE>(fa<SAT>(u1, u3), kn);

    o[0].re = fa<SAT>(s00, s02); o[0].im = fa<SAT>(s01, s03);
    o[1].re = fa<SAT>(s10, s12); o[1].im = fa<SAT>(s11, s13);
    o[2].re = fs<SAT>(s04, s06); o[2].im = fs<SAT>(s05, s07);
    o[3].re = fa<SAT>(s08, s14); o[3].im = fa<SAT>(s09, s15);
    o[4].re = fs<SAT>(s00, s02); o[4].im = fs<SAT>(s01, s03);
    o[5].re = fs<SAT>(s10, s12); o[5].im = fs<SAT>(s11, s13);
    o[6].re = fa<SAT>(s04, s06); o[6].im = fa<SAT>(s05, s07);
    o[7].re = fs<SAT>(s08, s14); o[7].im = fs<SAT>(s09, s15);
}

// 15 = 3 x 5 prime-factor core on scaled inputs, natural-order output (fx/fft.c:611-915)
template <bool SAT> LC3_HD void core15(const cpx *a, cpx *z)
{
    cpx y[3][5], in[5];
#pragma unroll
    for (int k = 0; k < 3; k++) {
#pragma unroll
        for (int j = 0; j < 5; j++) in[j] = a[(5 * k + 3 * j) % 15];
        dft5<false, SAT>(in, y[k]);
    }
#pragma unroll
    for (int j = 0; j < 5; j++)
        dft3<SAT>(y[0][j], y[1][j], y[2][j], z[(6 * j) % 15], z[(6 * j + 10) % 15], z[(6 * j + 5) % 15]);
}

// Leaf transforms.  LD(m) returns input m already shifted right by the leaf's scale (5);
// ST(m, c) stores output m.  All loads precede all stores, so LD/ST may alias.
template <bool SAT, class LD, class ST> LC3_HD void fft10(LD ld, ST st)
{ // fx/fft.c:433-596 : 2 x 5
    cpx in[5], y[2][5];
#pragma unroll
    for (int k = 0; k < 2; k++) {
#pragma unroll
        for (int j = 0; j < 5; j++) in[j] = ld((5 * k + 6 * j) % 10);
        dft5<false, SAT>(in, y[k]);
    }
#pragma unroll
    for (int j = 0; j < 5; j++) {
        cpx p = {fa<SAT>(y[0][j].re, y[1][j].re), fa<SAT>(y[0][j].im, y[1][j].im)};
        cpx m = {fs<SAT>(y[0][j].re, y[1][j].re), fs<SAT>(y[0][j].im, y[1][j].im)};
        st(2 * j, p);
        st((2 * j + 5) % 10, m);
    }
}
template <bool SAT, class LD, class ST> LC3_HD void fft15(LD ld, ST st)
{ // fx/fft.c:611-915
    cpx a[15], z[15];
#pragma unroll
    for (int i = 0; i < 15; i++) a[i] = ld(i);
    core15<SAT>(a, z);
#pragma unroll
    for (int i = 0; i < 15; i++) st(i, z[i]);
}
template <bool SAT, class LD, class ST> LC3_HD void fft20(LD ld, ST st)
{ // fx/fft.c:1395-1775 : 4 x 5
    cpx in[5], y[4][5], c[4], o[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
#pragma unroll
        for (int j = 0; j < 5; j++) in[j] = ld((5 * k + 16 * j) % 20);
        dft5<false, SAT>(in, y[k]);
    }
#pragma unroll
    for (int j = 0; j < 5; j++) {
#pragma unroll
        for (int k = 0; k < 4; k++) c[k] = y[k][j];
        dft4<SAT>(c, o);
#pragma unroll
        for (int m = 0; m < 4; m++) st((4 * j + 5 * m) % 20, o[m]);
    }
}
template <bool SAT, class LD, class ST> LC3_HD void fft30(LD ld, ST st)
{ // fx/fft.c:1791-2598 : 2 x 15
    cpx a[15], z[2][15];
#pragma unroll
    for (int h = 0; h < 2; h++) {
#pragma unroll
        for (int i = 0; i < 15; i++) a[i] = ld((16 * i + 15 * h) % 30);
        core15<SAT>(a, z[h]);
    }
#pragma unroll
    for (int i = 0; i < 15; i++) {
        const int c = (8 * i) % 15;
        i32 r1 = z[0][c].re >> 1, r2 = z[1][c].re >> 1;
        i32 r3 = z[0][c].im >> 1, r4 = z[1][c].im >> 1;
        cpx p = {fa<SAT>(r1, r2), fa<SAT>(r3, r4)}, m = {fs<SAT>(r1, r2), fs<SAT>(r3, r4)};
        if (i & 1) { st(15 + i, p); st(i, m); }
        else       { st(i, p); st(15 + i, m); }
    }
}
template <int LEAF, bool SAT, class LD, class ST> LC3_HD void fft_leaf(LD ld, ST st)
{
    if (LEAF == 10) fft10<SAT>(ld, st);
    else if (LEAF == 15) fft15<SAT>(ld, st);
    else if (LEAF == 20) fft20<SAT>(ld, st);
    else fft30<SAT>(ld, st);
}

// fftN2 (fx/fft.c:3364-3651).  x: in/out (interleaved re,im), w: scratch of 2*LEAF*DIM2 words.
template <int LEAF, int DIM2, int SC, int WOFF, bool SAT, class A, class B> LC3_HD void fft_n2(A x, B w, const i16 *W)
{
    for (int i = 0; i < DIM2; i++) {
        fft_leaf<LEAF, SAT>(
            [&](int m) { cpx c = {x[2 * (i + m * DIM2)] >> 5, x[2 * (i + m * DIM2) + 1] >> 5}; return c; },
            [&](int m, cpx c) { w[2 * (i * LEAF + m)] = c.re; w[2 * (i * LEAF + m) + 1] = c.im; });
    }
    constexpr int SH = DIM2 == 4 ? 2 : 1;    // SCALEFACTOR60 - SCALEFACTOR15 | 1
    constexpr int TWS = DIM2 == 4 ? 8 : SC;  // twiddle row stride (fx/fft.c:3440)
    for (int i = 0; i < LEAF; i++) {
        cpx v[DIM2], o[DIM2];
#pragma unroll
        for (int k = 0; k < DIM2; k++) {
            i32 a = w[2 * (i + k * LEAF)], b = w[2 * (i + k * LEAF) + 1];
            if (k == 0 || i == 0) {
                v[k].re = a >> SH;
                v[k].im = b >> SH;
            } else {
                i32 c = W[SC * i + TWS * k * LEAF - WOFF], d = W[SC * i + TWS * k * LEAF + 1 - WOFF];
                v[k].re = fs<SAT>(mul16<SAT>(a, c), mul16<SAT>(b, d)) >> SH;
                v[k].im = fa<SAT>(mul16<SAT>(a, d), mul16<SAT>(b, c)) >> SH;
            }
        }
        if (DIM2 == 4) dft4<SAT>(v, o); else dft8<false, SAT>(v, o);
#pragma unroll
        for (int m = 0; m < DIM2; m++) { x[2 * (i + m * LEAF)] = o[m].re; x[2 * (i + m * LEAF) + 1] = o[m].im; }
    }
}

// fft40 (fx/fft.c:3208-3362)
template <bool SAT, class A, class B> LC3_HD void fft40(A x, B w)
{
    for (int i = 0; i < 8; i++) {
        cpx in[5], o[5];
#pragma unroll
        for (int j = 0; j < 5; j++) { in[j].re = x[2 * (i + j * 8)] >> 4; in[j].im = x[2 * (i + j * 8) + 1] >> 4; }
        dft5<true, SAT>(in, o);
#pragma unroll
        for (int j = 0; j < 5; j++) { w[2 * (i * 5 + j)] = o[j].re; w[2 * (i * 5 + j) + 1] = o[j].im; }
    }
    for (int i = 0; i < 5; i++) {
        cpx v[8], o[8];
#pragma unroll
        for (int k = 0; k < 8; k++) {
            i32 a = w[2 * (i + k * 5)], b = w[2 * (i + k * 5) + 1];
            if (k == 0 || i == 0) {
                v[k].re = a >> 1;
                v[k].im = b >> 1;
            } else {
                i32 c = rom_rot40[2 * (i - 1) + (k - 1) * 8], d = rom_rot40[2 * (i - 1) + (k - 1) * 8 + 1];
                v[k].re = fa<SAT>(mpy32x32k(a, c), mpy32x32k(b, d)) >> 1;
                v[k].im = fs<SAT>(mpy32x32k(b, c), mpy32x32k(a, d)) >> 1;
            }
        }
        dft8<true, SAT>(v, o);
#pragma unroll
        for (int m = 0; m < 8; m++) { x[2 * (i + m * 5)] = o[m].re; x[2 * (i + m * 5) + 1] = o[m].im; }
    }
}

// BASOP_cfft dispatch (fx/fft.c:3668-3775); returns the scale added to the block exponent
template <int LEN, class A, class B> LC3_HD int cfft(A x, B w)
{
    constexpr bool SAT = FftNeedsSat<LEN>::v;
    auto ldx = [&](int m) { cpx c = {x[2 * m] >> 5, x[2 * m + 1] >> 5}; return c; };
    auto stx = [&](int m, cpx c) { x[2 * m] = c.re; x[2 * m + 1] = c.im; };
    if (LEN == 10) { fft10<SAT>(ldx, stx); return 5; }
    if (LEN == 20) { fft20<SAT>(ldx, stx); return 5; }
    if (LEN == 30) { fft30<SAT>(ldx, stx); return 6; }
    if (LEN == 40) { fft40<SAT>(x, w); return 7; }
    if (LEN == 60) { fft_n2<15, 4, 4, 60, SAT>(x, w, rom_rot480); return 7; }
    if (LEN == 80) { fft_n2<10, 8, 4, 40, SAT>(x, w, rom_rot320); return 8; }
    if (LEN == 120) { fft_n2<15, 8, 4, 60, SAT>(x, w, rom_rot480); return 8; }
    if (LEN == 160) { fft_n2<20, 8, 2, 40, SAT>(x, w, rom_rot320); return 8; }
    fft_n2<30, 8, 2, 60, SAT>(x, w, rom_rot480);
    return 9;
}

} // namespace lc3k
