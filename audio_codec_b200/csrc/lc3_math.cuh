// Table-driven fixed-point helpers of the encoder: log2, 2^x, 1/x, 1/sqrt, block headroom scans,
// pseudo-float add/compare.  Each follows fx/basop_util.c at the cited lines.
#pragma once
#include "lc3_rom.cuh"

namespace lc3k {

// BASOP_Util_Log2 (fx/basop_util.c:25-86): log2(x)/64 in Q31
// the mantissa polynomial of BASOP_Util_Log2 for f = 0 .. 16383 (the only values the normalised argument yields).
// tests/hostsim/primcheck.cpp runs this and the saturating operator-by-operator form over all 16384 values of f: they
// agree, i.e. none of the reference's L_mult / L_mac / mult saturations can trigger, and plain arithmetic is exact.
LC3_HD i32 m_log2_poly(i32 f)
{
    i32 acc = 2 * f * rom_ld_coeff[0];
    i32 p = (f * f) >> 15;
#pragma unroll
    for (int k = 1; k < 7; k++) {
        acc += 2 * p * rom_ld_coeff[k];
        if (k < 6) p = (p * f) >> 15;
    }
    acc = (acc >> 1) + hi16(acc) * 14506;
    return acc >> 5;
}
LC3_HD i32 m_log2(i32 x)
{
    if (x == 0) return LC3_I32_MIN;
    i32 e = l_norm(x);
    x = (i32)((u32)x << e);
    return l_sub(m_log2_poly(hi16(l_sub(LC3_I32_MAX, x))), e << 25);
}

// BASOP_Util_InvLog2 (fx/basop_util.c:89-138): 2^(x/64)
LC3_HD i32 m_invlog2(i32 x)
{
    if (x < -1040187392) return 0;
    if (x >= 1040187392 || x == 0) return LC3_I32_MAX;
    i32 frac = x & 0x3FF;
    u32 i3 = (u32)(x >> 10) & 0x1F, i2 = (u32)(x >> 15) & 0x1F, i1 = (u32)(x >> 20) & 0x1F;
    i32 e = lo16(x >> 25);
    if (x > 0) e = s_sub(31, e);
    if (x < 0) e = s_neg(e);
    u32 l3 = (u32)l_add((i32)rom_exp2x_tab[i3], mpy32x16k(0x0016302F, frac) >> 1);
    u32 l12 = (u32)mpy32x32((i32)rom_exp2_tab[i1], (i32)rom_exp2w_tab[i2]);
    u32 l = (u32)mpy32x32((i32)l12, (i32)l3);
    return l_shr((i32)l, s_sub(e, 3));
}

// ISqrt16 (fx/basop_util.c:165-185, :212-234); *e in/out
LC3_HD i32 m_isqrt16(i32 m, i32 *e)
{
    i32 pre = s_norm(m);
    i32 ee = s_sub(*e, pre);
    m = s_shl(m, pre);
    i32 idx = s_mac_r(-32768 - (32 << 16), m, 1 << 6);
    i32 fr = m & 0x1FF;
    m = s_msu_r(rom_isqrt16_tab[idx], rom_isqrt16_diff[idx], fr);
    if ((ee & 1) == 0) m = s_mult_r(m, 0x5a82);
    *e = s_msu_r((i32)1 << 15, ee, 1 << 14);
    return m;
}

// Inv16 (fx/basop_util.c:236-273); *e in/out
LC3_HD i32 m_inv16(i32 mant, i32 *e)
{
    i32 m = s_abs(s_max(mant, -32767));
    i32 pre = s_norm(m);
    i32 ee = s_sub(*e, pre);
    m = s_shl(m, pre);
    i32 idx = s_mac_r(-32768 - (32 << 16), m, 1 << 6);
    i32 fr = s_shl(m & 0x1FF, 1);
    m = s_msu_r(rom_inv16_tab[idx], rom_inv16_diff[idx], fr);
    if (mant < 0) m = s_neg(m);
    *e = s_sub(1, ee);
    return m;
}

// Isqrt (fx/basop_util.c:788-832)
LC3_HD i32 m_isqrt32(i32 x, i32 *e)
{
    if (x <= 0) { *e = 0; return LC3_I32_MAX; }
    i32 odd = *e & 1;
    i32 idx = (x >> 23) & 0x7f;
    i32 fr = hi16((i32)((u32)(x & 0x007FFFFF) << 8));
    i32 d = l_sub(rom_isqrt32_tab[idx + 1], rom_isqrt32_tab[idx]);
    i32 y = l_add(rom_isqrt32_tab[idx], mpy32x16(d, fr));
    if (odd) { y = mpy32x16k(y, 0x5A82); y = l_shl(y, odd); }
    *e = s_mac_r(32768, *e, -16384);
    return y;
}

// BASOP_Util_Log2_16 (fx/basop_util.c:835-861)
LC3_HD i32 m_log2_16(i32 x, i32 x_e)
{
    if (x == 0) return -32768;
    i32 sh = l_norm(x);
    x = l_shl(x, s_sub(sh, 10));
    i32 ip = s_shl(s_sub(s_sub(x_e, sh), 1), 9);
    i32 t1 = s_mac_r(x, -33, 16384);
    i32 t2 = s_lshr(lo16(x), 6);
    i32 fp = s_mac_r(rom_log2_16_t1[t1], rom_log2_16_t2[t1], t2);
    return s_add(ip, fp);
}

// BASOP_Util_InvLog2_16 (fx/basop_util.c:863-873)
LC3_HD i32 m_invlog2_16(i32 x, i32 *y_e)
{
    i32 t1 = (x & 2047) >> 5;
    i32 t2 = (x & 31) << 4;
    i32 y = s_mac_r(rom_invlog2_16_t1[t1], rom_invlog2_16_t2[t1], t2);
    *y_e = s_add(x >> 11, 1);
    return y;
}

// Headroom of a block from its running positive maximum / negative minimum
// (getScaleFactor16/_0 fx/basop_util.c:286-357, getScaleFactor32/_0 :368-444).
struct Range { i32 mx, mn; };
LC3_HD void range_init(Range &r) { r.mx = 0; r.mn = 0; }
LC3_HD void range_add(Range &r, i32 v) { r.mx = imax(r.mx, v); r.mn = imin(r.mn, v); }
LC3_HD i32 range_headroom16(const Range &r, int zero_val)
{
    i32 a = 16, b = 16;
    if (r.mx != 0) a = s_norm(r.mx);
    if (r.mn != 0) b = s_norm(r.mn);
    i32 h = imin(a, b);
    return zero_val ? h : (h & 0xF);
}
LC3_HD i32 range_headroom32(const Range &r, int zero_val)
{
    i32 a = 32, b = 32;
    if (r.mx != 0) a = l_norm(r.mx);
    if (r.mn != 0) b = l_norm(r.mn);
    i32 h = imin(a, b);
    return zero_val ? h : (h & 0x1F);
}
template <class A> LC3_HD i32 m_headroom16(A x, int n, int zero_val)
{
    Range r; range_init(r);
    for (int i = 0; i < n; i++) range_add(r, x[i]);
    return range_headroom16(r, zero_val);
}
template <class A> LC3_HD i32 m_headroom32(A x, int n, int zero_val)
{
    Range r; range_init(r);
    for (int i = 0; i < n; i++) range_add(r, x[i]);
    return range_headroom32(r, zero_val);
}

// BASOP_Util_Divide3216_Scale (fx/basop_util.c:446-482)
LC3_HD i32 m_div3216(i32 x, i32 y, i32 *s)
{
    if (x == 0) { *s = 0; return 0; }
    i32 sign = hi16(x) ^ y;
    x = l_abs(l_max(x, LC3_I32_MIN + 1));
    y = s_abs(s_max(y, -32767));
    i32 sx = s_sub(l_norm(x), 1);
    x = l_shl(x, sx);
    i32 sy = s_norm(y);
    y = s_shl(y, sy);
    *s = s_sub(sy, sx);
    i32 z = s_div(l_round(x), y);
    if (sign < 0) z = s_neg(z);
    return z;
}

// BASOP_Util_Divide1616_Scale (fx/basop_util.c:484-532)
LC3_HD i32 m_div1616(i32 x, i32 y, i32 *s)
{
    i32 sign = 0;
    if (x < 0) { x = s_neg(x); sign ^= 1; }
    if (y < 0) { y = s_neg(y); sign ^= 1; }
    if (x == 0) { *s = 0; return 0; }
    i32 sx = s_norm(x);
    x = s_shr(s_shl(x, sx), 1);
    *s = s_sub(1, sx);
    i32 sy = s_norm(y);
    y = s_shl(y, sy);
    *s = s_add(*s, sy);
    i32 z = s_div(x, y);
    if (sign) z = s_neg(z);
    return z;
}

// Scale_sig (fx/basop_util.c:571-595) on one value; block form below
LC3_HD i32 scale_val(i32 v, i32 e, i32 t) { return e > 0 ? s_shl(v, e) : s_msu_r(0, v, t); }
template <class A> LC3_HD void m_scale_sig(A x, int n, i32 e)
{
    if (e == 0) return;
    i32 t = e < 0 ? s_shl(-32768, s_max(e, -15)) : 0;
    for (int i = 0; i < n; i++) x[i] = (i16)scale_val(x[i], e, t);
}

// BASOP_Util_Add_Mant32Exp (fx/basop_util.c:633-684)
LC3_HD i32 m_add_me(i32 am, i32 ae, i32 bm, i32 be, i32 *re)
{
    if (!am) ae = be;
    if (!bm) be = ae;
    i32 sh = s_sub(ae, be);
    sh = imin(31, imax(-31, sh));
    if (sh < 0) am = l_shl(am, sh);
    if (sh > 0) bm = l_shr(bm, sh);
    ae = s_add(imax(ae, be), 1);
    i32 t = l_add(am >> 1, bm >> 1);
    sh = l_norm(t);
    if (sh) t = l_shl(t, sh);
    if (t == 0) ae = 0; else ae = s_sub(ae, sh);
    *re = ae;
    return t;
}

// BASOP_Util_Cmp_Mant32Exp (fx/basop_util.c:686-768): 1 if a > b, 0 if equal, -1 otherwise
LC3_HD i32 m_cmp_me(i32 am, i32 ae, i32 bm, i32 be)
{
    i32 sh = l_norm(am);
    if (sh) { am = l_shl(am, sh); ae = s_sub(ae, sh); }
    sh = l_norm(bm);
    if (sh) { bm = l_shl(bm, sh); be = s_sub(be, sh); }
    if (!am) ae = be;
    if (!bm) be = ae;
    i32 dm;
    if (am > 0 && bm < 0) dm = 1;
    else if (am < 0 && bm > 0) dm = -1;
    else dm = l_sub(am, bm);
    i32 de = s_sub(ae, be);
    if (dm == 0 && de == 0) return 0;
    i32 r = -1;
    if (am >= 0) {
        if (bm < 0) r = 1;
        if (bm >= 0 && (de > 0 || (de == 0 && dm > 0))) r = 1;
    } else {
        if (bm < 0 && (de < 0 || (de == 0 && dm > 0))) r = 1;
    }
    return r;
}

} // namespace lc3k
