// Low-delay MDCT of one frame per lane: window + time-domain aliasing fold (fx/mdct_fx.c:17-96),
// block normalisation, then DCT-IV = pre-twiddle -> complex FFT(N/2) -> post-twiddle
// (fx/dct4_fx.c:16-138; twiddle tables fx/rom_basop_util.c:590-670).
#pragma once
#include "lc3_fft.cuh"
#include "lc3_math.cuh"

namespace lc3k {

// (a + jb)(c + jd), every product halved before the add (cplxMpy32_32_16_2, fx/basop_mpy.h:45-50)
// The twiddle tables hold no -32768 (tests/test_rom_bounds.py), so each halved product is below 2^30 in
// magnitude: neither Mpy_32_16's -1 x -1 case nor the L_add/L_sub saturation of the reference can trigger.
LC3_HD void cmul_half(i32 &re, i32 &im, i32 a, i32 b, i32 c, i32 d)
{
    re = (mpy32x16k(a, c) >> 1) - (mpy32x16k(b, d) >> 1);
    im = (mpy32x16k(a, d) >> 1) + (mpy32x16k(b, c) >> 1);
}

template <int L> struct TwSlot {
    static constexpr int v = L == 20 ? 0 : L == 40 ? 1 : L == 60 ? 2 : L == 80 ? 3 : L == 120 ? 4 : L == 160 ? 5 : L == 240 ? 6 : L == 320 ? 7 : 8;
};

// DCT-IV of one frame in place (fx/dct4_fx.c:16-138): pre-twiddle -> complex FFT(N/2) -> post-twiddle.  The block shift s
// (the headroom of y: the shift cannot overflow) is applied as the pre-twiddle loads.  Returns the scale of the FFT; the
// caller adds it and then 2 to the block exponent (dct4_fx.c:93, :136).  Used by the encoder's MDCT below and by the
// decoder's inverse MDCT (fx/imdct_fx.c:84).
template <int N, class A, class B> LC3_HD i32 dct4_frame(A y, B w, i32 s)
{
    constexpr int M = N >> 1;
    const i16 *tw = &rom_dct4_twiddle[rom_dct4_twiddle_off[TwSlot<N>::v]];
    const i16 *sn = (N % 3 == 0) ? rom_sin480 : rom_sin320;
    constexpr int step = (N % 3 == 0) ? 960 / N : 640 / N;
    i32 a1, a2, a3, a4;
    for (int k = 0, lo = 0, hi = N - 2; k < M - 1; k += 2, lo += 2, hi -= 2) {
        // s is the block headroom: the shift cannot overflow
        i32 p0 = (i32)((u32)y[lo] << s), p1 = (i32)((u32)y[lo + 1] << s), q0 = (i32)((u32)y[hi] << s), q1 = (i32)((u32)y[hi + 1] << s);
        cmul_half(a1, a2, q1, p0, tw[2 * k], tw[2 * k + 1]);
        cmul_half(a3, a4, q0, p1, tw[2 * k + 2], tw[2 * k + 3]);
        y[lo] = a2; y[lo + 1] = a1;
        y[hi] = a4; y[hi + 1] = -a3; // cmul_half outputs are below 2^31 in magnitude: no L_negate saturation
    }
    const i32 fft_scale = cfft<M>(y, w);

    int lo = 0, hi = N - 2, idx = step;
    constexpr int M2 = ((M + 1) >> 1) - 1;
    cmul_half(a3, a4, y[hi], y[hi + 1], sn[2 * idx], sn[2 * idx + 1]);
    y[hi + 1] = -(y[lo + 1] >> 1);
    y[lo] = y[lo] >> 1;
    for (int k = 1; k < M2; k++) {
        y[lo + 1] = a3; y[hi] = a4;
        lo += 2; hi -= 2;
        cmul_half(a1, a2, y[lo + 1], y[lo], sn[2 * idx], sn[2 * idx + 1]);
        idx += step;
        cmul_half(a3, a4, y[hi], y[hi + 1], sn[2 * idx], sn[2 * idx + 1]);
        y[hi + 1] = -a1;
        y[lo] = a2;
    }
    y[lo + 1] = a3; y[hi] = a4;
    lo += 2; hi -= 2;
    cmul_half(a3, a4, y[lo + 1], y[lo], sn[2 * idx], sn[2 * idx + 1]);
    a1 = mpy32x16k(y[hi], 23170) >> 1;
    a2 = mpy32x16k(y[hi + 1], 23170) >> 1;
    y[hi] = a1 + a2; // |a1|, |a2| < 2^30
    y[lo + 1] = a1 - a2;
    y[hi + 1] = -a3;
    y[lo] = a4;
    return fft_scale;
}

// XS(i): scaled 16-bit input sample i of this frame, i in [-memLen, N) (negative = previous frame).
// y: N words, receives the spectrum.  w: N words of scratch.  Returns the block exponent.
template <int N, class XS, class A, class B>
LC3_HD i32 mdct_frame(XS xs, i32 x_exp, const i16 *win, int wLen, int memLen, A y, B w)
{
    constexpr int m = N >> 1;
    const int z = 2 * N - wLen;
    // |x| <= 2^15 and |w| <= 32767 (no -32768 in the window ROM): a sum of two products stays below 2^31, so the
    // reference's L_mac0/L_msu0 cannot saturate here
    Range rg; range_init(rg);
    int i;
    for (i = 0; i < m; i++) {
        i32 v = xs(i - memLen) * win[i] - xs(2 * m - 1 - i - memLen) * win[2 * m - 1 - i];
        y[m + i] = v; range_add(rg, v);
    }
    for (i = 0; i < z; i++) {
        i32 v = xs(2 * m - memLen + i) * win[2 * m + i];
        y[m - 1 - i] = v; range_add(rg, v);
    }
    for (; i < m; i++) {
        i32 v = xs(2 * m - memLen + i) * win[2 * m + i] + xs(4 * m - memLen - 1 - i) * win[4 * m - 1 - i];
        y[m - 1 - i] = v; range_add(rg, v);
    }
    const i32 s = imax(0, range_headroom32(rg, 0));
    i32 y_e = s_sub(s_sub(x_exp, 2), s);
    if (N <= 20) y_e = s_add(y_e, 2);
    else if (N <= 120) y_e = s_add(y_e, 1);

    y_e = s_add(y_e, dct4_frame<N>(y, w, s));
    return s_add(y_e, 2);
}

} // namespace lc3k
