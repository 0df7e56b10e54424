// Profiling aid, compiled in only with -DLC3_DBG_TIMING (make NVFLAGS+=...): every warp (or block) of a kernel
// records globaltimer marks and event counts into a per-translation-unit device buffer, which
// tools/warp_times.py reads back through lc3k_dbg_read_<unit>().  It answers what ncu's averages hide: which warps
// are the stragglers and in which section of the body (k_tns ran 2.5 ms because 2 % of the warps took 8x the
// median in the lattice filter and each held a one-block SM).  Row layout: LC3_DBG_COLS u64 per work unit;
// columns 0..5 time marks, 6..7 counters.  Nothing of this exists in a normal build.
#pragma once

#define LC3_DBG_ROWS 16384
#define LC3_DBG_COLS 8
#define LC3_DBG_KERNELS 4

#if defined(LC3_DBG_TIMING) && defined(__CUDACC__)
static __device__ unsigned long long lc3k_dbg[LC3_DBG_KERNELS][LC3_DBG_ROWS][LC3_DBG_COLS];
#define LC3_DBG_READER(unit) \
    extern "C" int lc3k_dbg_read_##unit(void *dst) { return (int)cudaMemcpyFromSymbol(dst, lc3k_dbg, sizeof(lc3k_dbg)); }
#else
#define LC3_DBG_READER(unit)
#endif

#if defined(LC3_DBG_TIMING) && defined(__CUDA_ARCH__)
// the row of the running warp is set by the kernel before it calls its body
#define LC3_DBG_ENTER(kernel, row) const int lc3_dbg_k_ = (kernel), lc3_dbg_r_ = (row); \
    if ((threadIdx.x & 31) == 0 && lc3_dbg_r_ < LC3_DBG_ROWS) { lc3k_dbg[lc3_dbg_k_][lc3_dbg_r_][6] = 0; lc3k_dbg[lc3_dbg_k_][lc3_dbg_r_][7] = 0; } \
    LC3_DBG_MARK_AT(lc3_dbg_k_, lc3_dbg_r_, 0)
#define LC3_DBG_LEAVE() LC3_DBG_MARK_AT(lc3_dbg_k_, lc3_dbg_r_, 5)
#define LC3_DBG_MARK_AT(kernel, row, col) do { __syncwarp(); unsigned long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); \
    if ((threadIdx.x & 31) == 0 && (row) < LC3_DBG_ROWS) lc3k_dbg[kernel][row][col] = t_; } while (0)
#define LC3_DBG_COUNT_AT(kernel, row, col) do { if ((row) < LC3_DBG_ROWS) atomicAdd(&lc3k_dbg[kernel][row][col], 1ull); } while (0)
#else
#define LC3_DBG_ENTER(kernel, row) ((void)0)
#define LC3_DBG_LEAVE() ((void)0)
#define LC3_DBG_MARK_AT(kernel, row, col) ((void)0)
#define LC3_DBG_COUNT_AT(kernel, row, col) ((void)0)
#endif
