// Saturating 16/32-bit fixed-point primitives for the sm_100a LC3plus encoder kernels.
//
// Semantics are those of the ITU-T STL2009 basic operators the ETSI reference is written in
// (fx/basic_op/basop32.c; Mpy_32_16_ss / Mpy_32_32_ss fx/basic_op/enh40.c:126-251;
// fx/basop_mpy.c:16-35).  Every operator is a pure function of its arguments (the reference build
// hides the Overflow/Carry globals, fx/basic_op/basop32.h:50-51), so one GPU thread can evaluate
// them independently.  Written so that each maps to a short SASS sequence: saturation through
// min/max or a sign test, normalisation through CLZ, fractional products through IMAD.WIDE + SHF.
//
// The file also compiles as plain C++ (LC3_HD empty): tests/hostsim builds the same stage code for
// the CPU to debug kernels where there is no GPU.  That build is test tooling; the product
// library contains the CUDA build only.
#pragma once
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#if defined(__CUDACC__)
#define LC3_HD __device__ __forceinline__
#define LC3_HDN __device__ __noinline__
#define LC3_HDM __device__ __forceinline__
#else
#define LC3_HD static inline
#define LC3_HDN static
#define LC3_HDM inline
#endif

namespace lc3k {

typedef int16_t i16;
typedef int32_t i32;
typedef uint32_t u32;
typedef int64_t i64;
typedef uint8_t u8;

#define LC3_I32_MAX 2147483647
#define LC3_I32_MIN (-2147483647 - 1)

LC3_HD i32 imin(i32 a, i32 b) { return a < b ? a : b; }
LC3_HD i32 imax(i32 a, i32 b) { return a > b ? a : b; }
// (With LC3_SAT16_PTX the clamp is written in PTX: left as C, the compiler recognises "saturating 16-bit add/sub" around
// values loaded from int16 arrays and lowers it to a dozen byte-permute / compare / xor instructions per call, which is
// what the SNS codebook search is made of.  The other kernels measure slightly faster with the C form.)
LC3_HD i32 sat16(i32 v)
{
#if defined(__CUDA_ARCH__) && defined(LC3_SAT16_PTX)
    i32 r;
    asm("min.s32 %0, %1, 32767;\n\tmax.s32 %0, %0, -32768;" : "=r"(r) : "r"(v));
    return r;
#else
    return imax(imin(v, 32767), -32768);
#endif
}
LC3_HD i32 sat32(i64 v) { return (i32)(v > LC3_I32_MAX ? LC3_I32_MAX : (v < LC3_I32_MIN ? LC3_I32_MIN : v)); }

// ---- 16-bit operators.  Values are carried in 32-bit registers holding a sign-extended int16.
LC3_HD i32 s_add(i32 a, i32 b) { return sat16(a + b); }
LC3_HD i32 s_sub(i32 a, i32 b) { return sat16(a - b); }
LC3_HD i32 s_abs(i32 a) { return imin(a < 0 ? -a : a, 32767); }
LC3_HD i32 s_neg(i32 a) { return imin(-a, 32767); }
LC3_HD i32 s_max(i32 a, i32 b) { return imax(a, b); }
LC3_HD i32 s_min(i32 a, i32 b) { return imin(a, b); }
LC3_HD i32 s_shr(i32 a, i32 n);
LC3_HD i32 s_shl(i32 a, i32 n)
{
    if (n < 0) return s_shr(a, n < -16 ? 16 : -n);
    if (n > 15) return a == 0 ? 0 : (a > 0 ? 32767 : -32768);
    return sat16(a * (1 << n));                  // |a| <= 2^15, n <= 15: no 32-bit overflow
}
LC3_HD i32 s_shr(i32 a, i32 n)
{
    if (n < 0) return s_shl(a, n < -16 ? 16 : -n);
    return a >> imin(n, 15);
}
LC3_HD i32 s_shr_r(i32 a, i32 n)
{
    if (n > 15) return 0;
    i32 r = s_shr(a, n);
    if (n > 0 && ((a >> (n - 1)) & 1)) r++;
    return r;
}
LC3_HD i32 s_mult(i32 a, i32 b) { return imin((a * b) >> 15, 32767); }
LC3_HD i32 s_mult_r(i32 a, i32 b) { return imin((a * b + 0x4000) >> 15, 32767); }
LC3_HD i32 s_imult(i32 a, i32 b) { return sat16(a * b); }
LC3_HD i32 clz32(u32 v)
{
#if defined(__CUDA_ARCH__)
    return __clz((int)v);
#else
    return v ? __builtin_clz(v) : 32;
#endif
}
LC3_HD i32 s_norm(i32 a) { return a == 0 ? 0 : clz32((u32)(a ^ (a >> 31))) - 17; }
LC3_HD i32 s_lshr(i32 a, i32 n) { return (i32)(i16)((a & 0xffff) >> n); }
LC3_HD i32 sx16(i32 a) { return (i32)(i16)a; }   // extract_l: keep the low 16 bits, sign-extended

// ---- 32-bit operators
// a + b where the caller has shown that the sum fits 32 bits.  The host build (tests/hostsim) checks the claim on
// every call and stops at the first sum that would have wrapped.
LC3_HD i32 add_proved(i32 a, i32 b)
{
#if !defined(__CUDA_ARCH__)
    if ((i64)a + b != (i64)(i32)((u32)a + (u32)b)) { fprintf(stderr, "add_proved: %d + %d leaves 32 bits\n", a, b); abort(); }
#endif
    return (i32)((u32)a + (u32)b);
}
LC3_HD i32 l_add(i32 a, i32 b)
{
#if defined(__CUDA_ARCH__)
    i32 r;
    asm("add.sat.s32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
#else
    return sat32((i64)a + b);
#endif
}
LC3_HD i32 l_sub(i32 a, i32 b)
{
#if defined(__CUDA_ARCH__)
    i32 r;
    asm("sub.sat.s32 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
    return r;
#else
    return sat32((i64)a - b);
#endif
}
LC3_HD i32 l_abs(i32 a) { return a == LC3_I32_MIN ? LC3_I32_MAX : (a < 0 ? -a : a); }
LC3_HD i32 l_neg(i32 a) { return a == LC3_I32_MIN ? LC3_I32_MAX : -a; }
LC3_HD i32 l_max(i32 a, i32 b) { return imax(a, b); }
LC3_HD i32 l_min(i32 a, i32 b) { return imin(a, b); }
// L_shr with a non-negative count (L_shr_pos and every call site with a literal count)
LC3_HD i32 l_shr_p(i32 a, i32 n) { return a >> imin(n, 31); }
// L_shl with a non-negative count
LC3_HD i32 l_shl_p(i32 a, i32 n)
{
    if (n > 31) return a == 0 ? 0 : (a > 0 ? LC3_I32_MAX : LC3_I32_MIN);
    i32 r = (i32)((u32)a << n);
    return (r >> n) != a ? ((a >> 31) ^ LC3_I32_MAX) : r;
}
// L_shl for a count known not to exceed the headroom of a (norm_l(a) >= n): no saturation test
LC3_HD i32 l_shl_ns(i32 a, i32 n) { return n >= 0 ? (i32)((u32)a << n) : a >> imin(-n, 31); }
LC3_HD i32 l_shl(i32 a, i32 n) { return n <= 0 ? l_shr_p(a, n < -32 ? 32 : -n) : l_shl_p(a, n); }
LC3_HD i32 l_shr(i32 a, i32 n) { return n < 0 ? l_shl_p(a, n < -32 ? 32 : -n) : l_shr_p(a, n); }
// L_shl by a count that stays the same over a loop: the count is split once into a left and a right part (one of them
// zero), and each element costs a shift, a check shift and a select, without branches.  Equals l_shl(a, n) for every
// a and n (tests/hostsim/primcheck.cpp); counts above 31 behave like 31 because only 0 and -1 survive a left shift by
// 31 unchanged, and those two give the same result for any larger count.
struct ShlPlan { int sl, sr; };
LC3_HD ShlPlan shl_plan(i32 n) { ShlPlan p = {imin(imax(n, 0), 31), imin(imax(-n, 0), 31)}; return p; }
LC3_HD i32 l_shl_plan(i32 a, const ShlPlan &p)
{
    const i32 l = (i32)((u32)a << p.sl);
    return (l >> p.sl) != a ? ((a >> 31) ^ LC3_I32_MAX) : (l >> p.sr);
}
LC3_HD i32 l_norm(i32 a) { return a == 0 ? 0 : clz32((u32)(a ^ (a >> 31))) - 1; }
LC3_HD i32 ul_norm(u32 a) { return a == 0 ? 0 : clz32(a); }
LC3_HD i32 l_mult0(i32 a, i32 b) { return a * b; }
LC3_HD i32 l_mult(i32 a, i32 b)
{
    i32 p = a * b;
    return p == 0x40000000 ? LC3_I32_MAX : p * 2;
}
LC3_HD i32 l_mac(i32 acc, i32 a, i32 b) { return l_add(acc, l_mult(a, b)); }
LC3_HD i32 l_msu(i32 acc, i32 a, i32 b) { return l_sub(acc, l_mult(a, b)); }
LC3_HD i32 l_mac0(i32 acc, i32 a, i32 b) { return l_add(acc, a * b); }
LC3_HD i32 l_msu0(i32 acc, i32 a, i32 b) { return l_sub(acc, a * b); }
LC3_HD i32 hi16(i32 a) { return a >> 16; }
LC3_HD i32 lo16(i32 a) { return (i32)(i16)a; }
LC3_HD i32 l_dep_h(i32 a) { return (i32)((u32)a << 16); }
// round_fx = extract_h(L_add(a, 0x8000)): high half plus the rounding bit; the one saturating case (a >= 0x7FFF8000)
// is the only way to reach 0x8000.  Checked against the operator form over all 2^32 inputs (primcheck --full).
LC3_HD i32 l_round(i32 a) { return imin((a >> 16) + ((a >> 15) & 1), 32767); }
LC3_HD i32 s_mac_r(i32 acc, i32 a, i32 b) { return l_round(l_mac(acc, a, b)); }
LC3_HD i32 s_msu_r(i32 acc, i32 a, i32 b) { return l_round(l_msu(acc, a, b)); }

// (x*y)>>15 and (x*y)>>31; the single -1 x -1 case saturates (enh40.c:131-134, :209-212)
LC3_HD i32 mpy32x16(i32 x, i32 y)
{
    if (x == LC3_I32_MIN && y == -32768) return LC3_I32_MAX;
    return (i32)(((i64)x * y) >> 15);
}
// same, for a 16-bit factor known to differ from -32768 (ROM twiddles and literal constants)
LC3_HD i32 mpy32x16k(i32 x, i32 y) { return (i32)(((i64)x * y) >> 15); }
LC3_HD i32 mpy32x32(i32 x, i32 y)
{
    if (x == LC3_I32_MIN && y == LC3_I32_MIN) return LC3_I32_MAX;
    return (i32)(((i64)x * y) >> 31);
}
LC3_HD i32 mpy32x32k(i32 x, i32 y) { return (i32)(((i64)x * y) >> 31); }

// div_s (basop32.c:1859+) for 0 <= num <= den, den > 0: 15-bit restoring division == floor(num*2^15/den)
LC3_HD i32 s_div(i32 num, i32 den)
{
    if (num == 0) return 0;
    if (num == den) return 32767;
    return (i32)(((u32)num << 15) / (u32)den);
}
// div_l (basop32.c:2036+): 32-bit numerator / 16-bit denominator, both positive
LC3_HD i32 l_div(i32 num, i32 den)
{
    i32 d = l_dep_h(den);
    if (num >= d) return 32767;
    num >>= 1;
    d >>= 1;
    // restoring division on (num<<k) >= d: equals floor(num*2^15/d) while num < d
    return (i32)((((uint64_t)(u32)num) << 15) / (u32)d);
}

} // namespace lc3k
