// Warp-cooperative helpers for the per-stream passes: ONE WARP works on one channel-stream and its 32
// lanes split the inner loops (filter taps, correlation lags, spectral bins).
//
// Code is written in phases.  Inside  LC3_LANES(l) ... LC3_LANES_END  the statements are per-lane work
// (lane index l); everything outside is warp-uniform: every lane computes the same scalar in its own
// registers.  Lanes talk to each other only through the warp's shared-memory block, written inside a
// LANES phase (each element by exactly one lane) and read after the phase's closing __syncwarp, and
// through the reductions below, which read the calling lane's own slot of a 32-entry scratch array.
// On the host (tests/hostsim) a LANES phase is a plain loop over 32 lanes and the reductions are loops, so
// the same source runs without a GPU.
#pragma once
#include "lc3_basop.cuh"

namespace lc3k {

#if defined(__CUDA_ARCH__)
#define LC3_LANES(l) { const int l = (int)(threadIdx.x & 31);
#define LC3_LANES_END } __syncwarp();
#define LC3_LANE0 if ((threadIdx.x & 31) == 0) {
#define LC3_LANE0_END } __syncwarp();
#define LC3_SYNC __syncwarp()
// a small per-lane array that lives across LANES phases: registers on the device, [n][32] on the host
#define LC3_PERLANE(T, name, n) T name[n]
#define LC3_PL(name, k, l) name[k]
LC3_HD i32 warp_max(const i32 *red) { return __reduce_max_sync(0xffffffffu, red[threadIdx.x & 31]); }
LC3_HD i32 warp_min(const i32 *red) { return __reduce_min_sync(0xffffffffu, red[threadIdx.x & 31]); }
LC3_HD i32 warp_sum(const i32 *red) { return __reduce_add_sync(0xffffffffu, red[threadIdx.x & 31]); }
// 64-bit reductions from the 32-bit warp-reduce instruction (one REDUX each instead of five shuffle/add rounds):
// a sum as three partial sums of 16 + 16 + 32 bits (32 terms below 2^16 cannot carry out of 32 bits; the top part wraps
// like the 64-bit sum does), a maximum as the maximum of the high words and then of the low words among its holders
LC3_HD i64 warp_sum64(const i64 *red)
{
    const i64 v = red[threadIdx.x & 31];
    const u32 lo = __reduce_add_sync(0xffffffffu, (u32)v & 0xffffu);
    const u32 mid = __reduce_add_sync(0xffffffffu, (u32)(v >> 16) & 0xffffu);
    const i32 hi = __reduce_add_sync(0xffffffffu, (i32)(v >> 32));
    return (i64)(((uint64_t)(u32)hi << 32) + ((uint64_t)mid << 16) + lo);
}
LC3_HD i64 warp_sum_wide(const i32 *red) // 64-bit sum of non-negative 32-bit terms
{
    const u32 v = (u32)red[threadIdx.x & 31];
    return (i64)(((uint64_t)__reduce_add_sync(0xffffffffu, v >> 16) << 16) + __reduce_add_sync(0xffffffffu, v & 0xffffu));
}
LC3_HD i64 warp_max64(const i64 *red)
{
    const i64 v = red[threadIdx.x & 31];
    const i32 h = (i32)(v >> 32);
    const i32 mh = __reduce_max_sync(0xffffffffu, h);
    const u32 ml = __reduce_max_sync(0xffffffffu, h == mh ? (u32)v : 0u);
    return (i64)(((uint64_t)(u32)mh << 32) | ml);
}
#else
#define LC3_LANES(l) for (int l = 0; l < 32; l++) {
#define LC3_LANES_END }
#define LC3_LANE0 {
#define LC3_LANE0_END }
#define LC3_SYNC ((void)0)
#define LC3_PERLANE(T, name, n) T name[n][32]
#define LC3_PL(name, k, l) name[k][l]
LC3_HD i32 warp_max(const i32 *red) { i32 v = red[0]; for (int i = 1; i < 32; i++) v = red[i] > v ? red[i] : v; return v; }
LC3_HD i32 warp_min(const i32 *red) { i32 v = red[0]; for (int i = 1; i < 32; i++) v = red[i] < v ? red[i] : v; return v; }
LC3_HD i32 warp_sum(const i32 *red) { i32 v = 0; for (int i = 0; i < 32; i++) v += red[i]; return v; }
LC3_HD i64 warp_sum64(const i64 *red) { i64 v = 0; for (int i = 0; i < 32; i++) v += red[i]; return v; }
LC3_HD i64 warp_sum_wide(const i32 *red) { i64 v = 0; for (int i = 0; i < 32; i++) v += red[i]; return v; }
LC3_HD i64 warp_max64(const i64 *red) { i64 v = red[0]; for (int i = 1; i < 32; i++) v = red[i] > v ? red[i] : v; return v; }
#endif

// scratch for the reductions, part of every warp's shared block
struct WarpRed { i64 q[32], q2[32], q3[32]; i32 a[32], b[32], c[32], d[32]; };

// saturate a 64-bit total to 32 bits: equals a chain of saturating L_add/L_mac0 when all terms have the
// same sign (monotone partial sums), or when no partial sum can leave the 32-bit range
LC3_HD i32 clamp64(i64 v) { return (i32)(v > LC3_I32_MAX ? LC3_I32_MAX : (v < LC3_I32_MIN ? LC3_I32_MIN : v)); }

// key for "maximum value, smallest index on ties"
LC3_HD i64 argmax_key(i32 value, int index) { return ((i64)value << 32) | (u32)(0x7fffffff - index); }
LC3_HD int argmax_index(i64 key) { return 0x7fffffff - (int)(u32)(key & 0xffffffff); }

} // namespace lc3k
