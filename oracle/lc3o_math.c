/* ORACLE (test infrastructure, not product code).
 * Table-driven fixed-point math helpers of the encoder: log2 / 2^x / 1/x / 1/sqrt, block headroom
 * scans, pseudo-float add/compare.  Follows fx/basop_util.c (line ranges in lc3o_internal.h). */
#include "lc3o_internal.h"

/* fx/basop_util.c:25-86 : log2(x)/64 in Q31 by a 7-term series of (1 - mantissa) */
i32 o_log2(i32 x)
{
    if (x == 0) return I32_MIN_;
    i16 e = l_norm(x);
    x = l_shl(x, e);
    i16 f = hi16(l_sub(I32_MAX_, x));
    i32 acc = l_mult(f, rom_ld_coeff[0]);
    i16 p = s_mult(f, f);
    for (int k = 1; k < 7; k++) {
        acc = l_mac(acc, p, rom_ld_coeff[k]);
        if (k < 6) p = s_mult(p, f);
    }
    acc = l_mac0(l_shr(acc, 1), hi16(acc), 14506);
    acc = l_shr(acc, 5);
    return l_sub(acc, l_shl((i32)e, 25));
}

/* fx/basop_util.c:89-138 : 2^(x/64), three table lookups and one linear term */
i32 o_invlog2(i32 x)
{
    if (x < -1040187392) return 0;
    if (l_sub(x, 1040187392) >= 0 || x == 0) return I32_MAX_;
    i16 frac = lo16(x & 0x3FF);
    u32 i3 = (u32)(l_shr(x, 10) & 0x1F), i2 = (u32)(l_shr(x, 15) & 0x1F), i1 = (u32)(l_shr(x, 20) & 0x1F);
    i16 e = lo16(l_shr(x, 25));
    if (x > 0) e = s_sub(31, e);
    if (x < 0) e = s_neg(e);
    u32 l3 = (u32)l_add((i32)rom_exp2x_tab[i3], l_shr(mpy32x16(0x0016302F, frac), 1));
    u32 l12 = (u32)mpy32x32((i32)rom_exp2_tab[i1], (i32)rom_exp2w_tab[i2]);
    u32 l = (u32)mpy32x32((i32)l12, (i32)l3);
    return l_shr((i32)l, s_sub(e, 3));
}

/* fx/basop_util.c:165-185, :212-234 */
i16 o_isqrt16(i16 m, i16 *e)
{
    i16 pre = s_norm(m);
    i16 ee = s_sub(*e, pre);
    m = s_shl(m, pre);
    i16 idx = s_mac_r(-32768 - (32 << 16), m, 1 << 6);
    i16 fr = (i16)(m & 0x1FF);
    m = s_msu_r(rom_isqrt16_tab[idx], rom_isqrt16_diff[idx], fr);
    if ((ee & 1) == 0) m = s_mult_r(m, 0x5a82);
    *e = s_msu_r((i32)1 << 15, ee, 1 << 14);
    return m;
}

/* fx/basop_util.c:236-273 */
i16 o_inv16(i16 mant, i16 *e)
{
    i16 m = s_abs(s_max(mant, I16_MIN_ + 1));
    i16 pre = s_norm(m);
    i16 ee = s_sub(*e, pre);
    m = s_shl(m, pre);
    i16 idx = s_mac_r(-32768 - (32 << 16), m, 1 << 6);
    i16 fr = s_shl((i16)(m & 0x1FF), 1);
    m = s_msu_r(rom_inv16_tab[idx], rom_inv16_diff[idx], fr);
    if (mant < 0) m = s_neg(m);
    *e = s_sub(1, ee);
    return m;
}

/* fx/basop_util.c:788-832 */
i32 o_isqrt32(i32 x, i16 *e)
{
    if (x <= 0) { *e = 0; return I32_MAX_; }
    i16 odd = (i16)(*e & 1);
    i32 idx = l_shr(x, 23) & 0x7f;
    i16 fr = hi16(l_shl(x & 0x007FFFFF, 8));
    i32 d = l_sub(rom_isqrt32_tab[idx + 1], rom_isqrt32_tab[idx]);
    i32 y = l_add(rom_isqrt32_tab[idx], mpy32x16(d, fr));
    if (odd) { y = mpy32x16(y, 0x5A82); y = l_shl(y, odd); }
    *e = s_mac_r(32768, *e, -16384);
    return y;
}

/* fx/basop_util.c:835-861 */
i16 o_log2_16(i32 x, i16 x_e)
{
    if (x == 0) return I16_MIN_;
    i16 sh = l_norm(x);
    x = l_shl(x, s_sub(sh, 10));
    i16 ip = s_shl(s_sub(s_sub(x_e, sh), 1), 9);
    i16 t1 = s_mac_r(x, -33, 16384);
    i16 t2 = s_lshr(lo16(x), 6);
    i16 fp = s_mac_r(rom_log2_16_t1[t1], rom_log2_16_t2[t1], t2);
    return s_add(ip, fp);
}

/* fx/basop_util.c:863-873 */
i16 o_invlog2_16(i16 x, i16 *y_e)
{
    i16 t1 = s_shr((i16)(x & 2047), 5);
    i16 t2 = s_shl((i16)(x & 31), 4);
    i16 y = s_mac_r(rom_invlog2_16_t1[t1], rom_invlog2_16_t2[t1], t2);
    *y_e = s_add(s_shr(x, 11), 1);
    return y;
}

/* fx/basop_util.c:286-357 : zero_val = 0 -> getScaleFactor16 (masked), 16 -> getScaleFactor16_0 */
i16 o_headroom16(const i16 *x, int n, int zero_val)
{
    i16 mx = 0, mn = 0;
    for (int i = 0; i < n; i++) {
        if (x[i] >= 0) mx = s_max(mx, x[i]);
        if (x[i] < 0) mn = s_min(mn, x[i]);
    }
    i16 a = 16, b = 16;
    if (mx != 0) a = s_norm(mx);
    if (mn != 0) b = s_norm(mn);
    i16 r = s_min(a, b);
    return zero_val ? r : (i16)(r & 0xF);
}

/* fx/basop_util.c:368-444 */
i16 o_headroom32(const i32 *x, int n, int zero_val)
{
    i32 mx = 0, mn = 0;
    for (int i = 0; i < n; i++) {
        if (x[i] >= 0) mx = l_max(mx, x[i]);
        if (x[i] < 0) mn = l_min(mn, x[i]);
    }
    i16 a = 32, b = 32;
    if (mx != 0) a = l_norm(mx);
    if (mn != 0) b = l_norm(mn);
    i16 r = s_min(a, b);
    return zero_val ? r : (i16)(r & 0x1F);
}

/* fx/basop_util.c:446-482 */
i16 o_div3216(i32 x, i16 y, i16 *s)
{
    if (x == 0) { *s = 0; return 0; }
    i16 sign = (i16)(hi16(x) ^ y);
    x = l_abs(l_max(x, I32_MIN_ + 1));
    y = s_abs(s_max(y, I16_MIN_ + 1));
    i16 sx = s_sub(l_norm(x), 1);
    x = l_shl(x, sx);
    i16 sy = s_norm(y);
    y = s_shl(y, sy);
    *s = s_sub(sy, sx);
    i16 z = s_div(l_round(x), y);
    if (sign < 0) z = s_neg(z);
    return z;
}

/* fx/basop_util.c:484-532 */
i16 o_div1616(i16 x, i16 y, i16 *s)
{
    i16 sign = 0;
    if (x < 0) { x = s_neg(x); sign ^= 1; }
    if (y < 0) { y = s_neg(y); sign ^= 1; }
    if (x == 0) { *s = 0; return 0; }
    i16 sx = s_norm(x);
    x = s_shr(s_shl(x, sx), 1);
    *s = s_sub(1, sx);
    i16 sy = s_norm(y);
    y = s_shl(y, sy);
    *s = s_add(*s, sy);
    i16 z = s_div(x, y);
    if (sign) z = s_neg(z);
    return z;
}

/* fx/basop_util.c:534-569 : energy with a sequentially adapted pre-shift */
i32 o_norm32norm(const i32 *x, i16 headroom, i16 n, i16 *res_e)
{
    i16 sh = headroom;
    i32 acc = 0;
    for (int i = 0; i < n; i++) {
        if (l_sub(acc, 0x40000000) >= 0) { sh = s_sub(sh, 1); acc = l_shr(acc, 2); }
        i16 t = l_round(l_shl(x[i], sh));
        acc = l_mac0(acc, t, t);
    }
    if (acc != 0) {
        i16 s = s_norm(n);
        i32 inc = l_shl(mpy32x16(0x0000fffd, s_shl(n, s)), s_sub(15, s));
        acc = l_add(acc, inc);
    }
    *res_e = s_sub(1, s_shl(sh, 1));
    return acc;
}

/* fx/basop_util.c:571-595 */
void o_scale_sig(i16 *x, int n, i16 e)
{
    if (e > 0) {
        for (int i = 0; i < n; i++) x[i] = s_shl(x[i], e);
    } else if (e < 0) {
        i16 t = s_shl(-32768, s_max(e, -15));
        for (int i = 0; i < n; i++) x[i] = s_msu_r(0, x[i], t);
    }
}

/* fx/basop_util.c:633-684 */
i32 o_add_me(i32 am, i16 ae, i32 bm, i16 be, i16 *re)
{
    if (!am) ae = be;
    if (!bm) be = ae;
    i16 sh = s_sub(ae, be);
    sh = s_min(31, s_max(-31, sh));
    if (sh < 0) am = l_shl(am, sh);
    if (sh > 0) bm = l_shr(bm, sh);
    ae = s_add(s_max(ae, be), 1);
    i32 t = l_add(l_shr(am, 1), l_shr(bm, 1));
    sh = l_norm(t);
    if (sh) t = l_shl(t, sh);
    if (t == 0) ae = 0; else ae = s_sub(ae, sh);
    *re = ae;
    return t;
}

/* fx/basop_util.c:686-768 : 1 if a > b, 0 if equal, -1 otherwise */
i16 o_cmp_me(i32 am, i16 ae, i32 bm, i16 be)
{
    i16 sh = l_norm(am);
    if (sh) { am = l_shl(am, sh); ae = s_sub(ae, sh); }
    sh = l_norm(bm);
    if (sh) { bm = l_shl(bm, sh); be = s_sub(be, sh); }
    if (!am) ae = be;
    if (!bm) be = ae;
    i32 dm;
    if (am > 0 && bm < 0) dm = 1;
    else if (am < 0 && bm > 0) dm = -1;
    else dm = l_sub(am, bm);
    i16 de = s_sub(ae, be);
    if (dm == 0 && de == 0) return 0;
    i16 r = -1;
    if (am >= 0) {
        if (bm < 0) r = 1;
        if (bm >= 0 && (de > 0 || (de == 0 && dm > 0))) r = 1;
    } else {
        if (bm < 0 && (de < 0 || (de == 0 && dm > 0))) r = 1;
    }
    return r;
}
