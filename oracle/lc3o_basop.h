/* ORACLE (test infrastructure, not product code).
 *
 * Saturating 16/32-bit fixed-point primitives with the exact semantics of the ITU-T STL2009
 * basic operators the ETSI LC3plus fixed-point encoder is written in
 * (fx/basic_op/basop32.c, enh40.c:126-251 Mpy_32_16_ss/Mpy_32_32_ss, enhUL32.c, fx/basop_mpy.c:16-35).
 * Restated as pure inline functions on int16_t/int32_t; pinned bit-for-bit against the compiled
 * reference operators in tests/test_oracle_basop.py.
 */
#ifndef LC3O_BASOP_H
#define LC3O_BASOP_H
#include <stdint.h>

typedef int16_t i16;
typedef int32_t i32;
typedef uint32_t u32;
typedef int64_t i64;

#define I16_MAX_ 32767
#define I16_MIN_ (-32768)
#define I32_MAX_ 2147483647
#define I32_MIN_ (-2147483647 - 1)

static inline i16 sat16(i32 v) { return (i16)(v > I16_MAX_ ? I16_MAX_ : (v < I16_MIN_ ? I16_MIN_ : v)); }
static inline i32 sat32(i64 v) { return (i32)(v > I32_MAX_ ? I32_MAX_ : (v < I32_MIN_ ? I32_MIN_ : v)); }

/* ---- 16-bit (basop32.c add/sub/abs_s/negate/shl/shr/mult/mult_r/shr_r/norm_s) */
static inline i16 s_add(i16 a, i16 b) { return sat16((i32)a + b); }
static inline i16 s_sub(i16 a, i16 b) { return sat16((i32)a - b); }
static inline i16 s_abs(i16 a) { return a == I16_MIN_ ? I16_MAX_ : (i16)(a < 0 ? -a : a); }
static inline i16 s_neg(i16 a) { return a == I16_MIN_ ? I16_MAX_ : (i16)-a; }
static inline i16 s_max(i16 a, i16 b) { return a > b ? a : b; }
static inline i16 s_min(i16 a, i16 b) { return a < b ? a : b; }
static inline i16 s_shr(i16 a, i16 n);
static inline i16 s_shl(i16 a, i16 n)
{
    if (n < 0) return s_shr(a, (i16)(n < -16 ? 16 : -n));
    if (a == 0) return 0;
    if (n > 15) return a > 0 ? I16_MAX_ : I16_MIN_;
    return sat16((i32)a * ((i32)1 << n));
}
static inline i16 s_shr(i16 a, i16 n)
{
    if (n < 0) return s_shl(a, (i16)(n < -16 ? 16 : -n));
    if (n >= 15) return (i16)(a < 0 ? -1 : 0);
    return (i16)(a >> n);
}
static inline i16 s_shr_r(i16 a, i16 n)
{
    if (n > 15) return 0;
    i16 r = s_shr(a, n);
    if (n > 0 && (a & ((i16)1 << (n - 1)))) r++;
    return r;
}
static inline i16 s_mult(i16 a, i16 b) { return sat16(((i32)a * b) >> 15); }
static inline i16 s_mult_r(i16 a, i16 b) { return sat16((((i32)a * b) + 0x4000) >> 15); }
static inline i16 s_imult(i16 a, i16 b) { return sat16((i32)a * b); }
static inline i16 s_norm(i16 a)
{
    if (a == 0) return 0;
    if (a == -1) return 15;
    if (a < 0) a = (i16)~a;
    i16 n = 0;
    while (a < 0x4000) { a = (i16)(a << 1); n++; }
    return n;
}
/* logical shifts on the 16-bit pattern (lshl/lshr) */
static inline i16 s_lshr(i16 a, i16 n) { return (i16)((uint16_t)a >> n); }
static inline i16 s_lshl(i16 a, i16 n) { return (i16)((uint16_t)a << n); }

/* ---- 32-bit (L_add/L_sub/L_abs/L_negate/L_shl/L_shr/norm_l/L_mult/L_mac/...) */
static inline i32 l_add(i32 a, i32 b) { return sat32((i64)a + b); }
static inline i32 l_sub(i32 a, i32 b) { return sat32((i64)a - b); }
static inline i32 l_abs(i32 a) { return a == I32_MIN_ ? I32_MAX_ : (a < 0 ? -a : a); }
static inline i32 l_neg(i32 a) { return a == I32_MIN_ ? I32_MAX_ : -a; }
static inline i32 l_max(i32 a, i32 b) { return a > b ? a : b; }
static inline i32 l_min(i32 a, i32 b) { return a < b ? a : b; }
static inline i32 l_shr(i32 a, i16 n);
static inline i32 l_shl(i32 a, i16 n)
{
    if (n <= 0) return l_shr(a, (i16)(n < -32 ? 32 : -n));
    if (a == 0) return 0;
    if (n > 31) return a > 0 ? I32_MAX_ : I32_MIN_;
    return sat32((i64)a * ((i64)1 << n));
}
static inline i32 l_shr(i32 a, i16 n)
{
    if (n < 0) return l_shl(a, (i16)(n < -32 ? 32 : -n));
    if (n >= 31) return a < 0 ? -1 : 0;
    return a >> n;
}
static inline i16 l_norm(i32 a)
{
    if (a == 0) return 0;
    if (a == -1) return 31;
    if (a < 0) a = ~a;
    i16 n = 0;
    while (a < 0x40000000) { a <<= 1; n++; }
    return n;
}
static inline i16 ul_norm(u32 a)
{
    if (a == 0) return 0;
    i16 n = 0;
    while (!(a & 0x80000000u)) { a <<= 1; n++; }
    return n;
}
static inline i32 l_mult(i16 a, i16 b)
{
    i32 p = (i32)a * b;
    return p == 0x40000000 ? I32_MAX_ : p * 2;
}
static inline i32 l_mult0(i16 a, i16 b) { return (i32)a * b; }
static inline i32 l_mac(i32 acc, i16 a, i16 b) { return l_add(acc, l_mult(a, b)); }
static inline i32 l_msu(i32 acc, i16 a, i16 b) { return l_sub(acc, l_mult(a, b)); }
static inline i32 l_mac0(i32 acc, i16 a, i16 b) { return l_add(acc, l_mult0(a, b)); }
static inline i32 l_msu0(i32 acc, i16 a, i16 b) { return l_sub(acc, l_mult0(a, b)); }
static inline i16 hi16(i32 a) { return (i16)(a >> 16); }
static inline i16 lo16(i32 a) { return (i16)a; }
static inline i32 l_dep_h(i16 a) { return (i32)a << 16; }
static inline i16 l_round(i32 a) { return hi16(l_add(a, 0x8000)); }
static inline i16 s_mac_r(i32 acc, i16 a, i16 b) { return l_round(l_mac(acc, a, b)); }
static inline i16 s_msu_r(i32 acc, i16 a, i16 b) { return l_round(l_msu(acc, a, b)); }

/* (x*y)>>15 and (x*y)>>31 with the single -1*-1 saturation case (enh40.c:126-251) */
static inline i32 mpy32x16(i32 x, i16 y)
{
    if (x == I32_MIN_ && y == I16_MIN_) return I32_MAX_;
    return (i32)(((i64)x * y) >> 15);
}
static inline i32 mpy32x32(i32 x, i32 y)
{
    if (x == I32_MIN_ && y == I32_MIN_) return I32_MAX_;
    return (i32)(((i64)x * y) >> 31);
}

/* fractional divides (basop32.c div_s :1859+, div_l :2036+); operands are in range at every
 * encoder call site, so the reference's abort() paths are not restated. */
static inline i16 s_div(i16 num, i16 den)
{
    if (num == 0) return 0;
    if (num == den) return I16_MAX_;
    i32 n = num, d = den;
    i16 q = 0;
    for (int k = 0; k < 15; k++) {
        q = (i16)(q << 1);
        n <<= 1;
        if (n >= d) { n -= d; q = (i16)(q + 1); }
    }
    return q;
}
static inline i16 l_div(i32 num, i16 den)
{
    i32 d = l_dep_h(den);
    if (num >= d) return I16_MAX_;
    num >>= 1;
    d >>= 1;
    i16 q = 0;
    for (int k = 0; k < 15; k++) {
        q = s_shl(q, 1);
        num = l_shl(num, 1);
        if (num >= d) { num = l_sub(num, d); q = s_add(q, 1); }
    }
    return q;
}
#endif
