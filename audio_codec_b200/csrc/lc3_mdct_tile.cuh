// Low-delay MDCT of one TILE (32 channel-frames) per thread block, spectrum resident in shared memory.
//
// Same arithmetic, operation for operation, as the lane-per-frame mdct_frame() (lc3_mdct.cuh; fx/mdct_fx.c:17-96,
// fx/dct4_fx.c:16-138, fftN2 fx/fft.c:3364-3651) -- only the work-to-thread mapping differs.  The lane-per-frame
// form makes ~9 passes over a 2 x N-word per-lane working set that lives in global/local memory; with ~75 000
// lanes resident that set is far larger than L2 and every pass goes to DRAM (ncu: 9.5 GB per launch at C1).
// Here a block of 256 threads owns the 32 frames of one tile, keeps them in shared memory and walks the phases
//   fold      thread = (tile lane, 1/8 of the samples): window + aliasing fold, coalesced reads of the staged PCM
//   twiddle   warp = frame, lane = pair of complex points: block shift + pre-twiddle
//   leaf      thread = (frame, leaf): 8 leaf FFTs of length N/16 per frame, in registers, written back in place
//   column    warp = frame, lane = column: radix-8 butterflies with twiddles (load all, sync, store all)
//   post      warp = frame, lane = pair: post-twiddle
//   store     thread = (tile lane, 1/8 of the bins): coalesced write into the tile-interleaved spectrum
// Frame lengths with an 8-column FFT only (N = 160, 240, 320, 480); shorter frames keep the lane-per-frame path.
//
// Shared-memory row of a frame: complex point j at word 2j + 2(j >> 3) (one complex of padding per 8, so that the
// 8 contiguous points a column task loads and the stride-8 points a leaf task loads are both conflict-free as
// 64-bit accesses); row stride S = 2 mod 4 words so that the four frames of a warp's leaf tasks fall in disjoint
// bank halves, and tile-lane-major access (fold, store) is at worst 2-way conflicted.
#pragma once
#include "lc3_mdct.cuh"
#include "lc3_stages.cuh"
#include "lc3_spec.cuh"
#include "lc3_warp.cuh"

namespace lc3k {

template <int N> struct MdctTile {
    static constexpr int M = N / 2, LEAF = M / 8;
    static constexpr int ROW = N + 2 * LEAF;
    static constexpr int S = ROW + ((6 - ROW % 4) % 4);
    static constexpr int THREADS = 256, PARTS = THREADS / 32;
    static constexpr int SC = (N % 3 == 0) ? (N == 480 ? 2 : 4) : (N == 320 ? 2 : 4); // twiddle stride of cfft<M> (lc3_fft.cuh)
    static constexpr int WOFF = (N % 3 == 0) ? 60 : 40;
    static constexpr int FFT_SCALE = (M == 240) ? 9 : 8;
    static constexpr bool SAT = FftNeedsSat<M>::v; // saturating FFT operators where the overflow bound does not close (lc3_fft.cuh)
    // twiddle tables, copied to shared memory by the kernel: in this mapping their indices differ from lane to lane,
    // which the constant cache serialises (one address per cycle); entries used: pre-twiddle N, post-twiddle
    // 2 * (M/2) * step + 2, column twiddles up to SC * (M - 1) + 2 - WOFF
    static constexpr int TW_LEN = N, SN_LEN = (N % 3 == 0) ? 482 : 322, W_LEN = 480;
    // per-band energies of the tile before alignment: mantissa and exponent [32 frames][BROW], row stride odd
    static constexpr int BROW = 65;
    static constexpr size_t smem_words = 32 * S + 3 * PARTS * 32 + 2 * 32 * BROW;
    static constexpr int WIN_LEN = 2 * N; // low-delay window (2N - la_zeroes entries), for the row-wise fold: lane-varying indices
    static constexpr size_t smem_bytes = 16 /* mbarrier of the row staging */ + smem_words * sizeof(i32) + (size_t)(TW_LEN + SN_LEN + W_LEN + WIN_LEN) * sizeof(i16);
    static_assert(N == 160 || N == 240 || N == 320 || N == 480, "8-column FFT lengths only");
};

struct MtTabs { const i16 *tw, *sn, *W; };
template <int N> LC3_HD MtTabs mt_rom_tabs()
{
    MtTabs T = {&rom_dct4_twiddle[rom_dct4_twiddle_off[TwSlot<N>::v]], (N % 3 == 0) ? rom_sin480 : rom_sin320,
                (N % 3 == 0) ? rom_rot480 : rom_rot320};
    return T;
}

LC3_HD int mt_ridx(int r) { return r + 2 * (r >> 4); }  // real index -> word in the padded row
LC3_HD int mt_cidx(int j) { return 2 * j + 2 * (j >> 3); } // complex index -> word of its real part

LC3_HD cpx mt_ld2(const i32 *p)
{
#if defined(__CUDA_ARCH__)
    const int2 v = *reinterpret_cast<const int2 *>(p);
    cpx c = {v.x, v.y};
#else
    cpx c = {p[0], p[1]};
#endif
    return c;
}
LC3_HD void mt_st2(i32 *p, cpx c)
{
#if defined(__CUDA_ARCH__)
    *reinterpret_cast<int2 *>(p) = make_int2(c.re, c.im);
#else
    p[0] = c.re; p[1] = c.im;
#endif
}

// block shift of frame fr from the per-part partial ranges
template <int N> LC3_HD i32 mt_shift(const i32 *mx, const i32 *mn, int fr)
{
    Range r; range_init(r);
    for (int p = 0; p < MdctTile<N>::PARTS; p++) { r.mx = imax(r.mx, mx[p * 32 + fr]); r.mn = imin(r.mn, mn[p * 32 + fr]); }
    return imax(0, range_headroom32(r, 0));
}

// ---- fold: thread t = (lane of the tile, part)
template <int N, int BITS> LC3_HD void mt_fold(const LC3K_Batch &b, int f, int tile, int t, i32 *X, i32 *mx, i32 *mn)
{
    typedef MdctTile<N> G;
    constexpr int m = N / 2;
    const int lane = t & 31, part = t >> 5, li = tile * LC3K_LANES + lane;
    i32 *y = X + lane * G::S;
    Range rg; range_init(rg);
    if (li < b.n_lanes) {
        // 16-bit input has the fixed exponent 15 (fx/enc_lc3.c:60-62), so this kernel does not wait for the pitch pass
        const i32 x_exp = BITS == 16 ? 15 : tile_view((const i32 *)b.pout, frame_tile(b, f, tile), PO_N, lane)[PO_X_EXP];
        PcmView<BITS> xs = pcm_view<BITS>(b, tile, lane, f, x_exp);
        const i16 *win = &rom_win[b.W_off];
        const int memLen = b.mdct_mem_len, z = 2 * N - b.W_len;
        // same three expressions as mdct_frame(); five outputs pairs at a time with all their samples fetched first
        // (twenty independent loads in flight per thread: this phase waits on global memory, not on arithmetic)
        constexpr int U = 5;
        static_assert((m / G::PARTS) % U == 0 && m % G::PARTS == 0, "fold unroll");
        for (int i0 = part; i0 < m; i0 += U * G::PARTS) {
            i32 xa[U], xb[U], xc[U], xd[U];
#pragma unroll
            for (int u = 0; u < U; u++) {
                const int i = i0 + u * G::PARTS;
                xa[u] = xs(i - memLen); xb[u] = xs(2 * m - 1 - i - memLen); xc[u] = xs(2 * m - memLen + i);
                xd[u] = i < z ? 0 : xs(4 * m - memLen - 1 - i); // the window is zero there (la_zeroes)
            }
#pragma unroll
            for (int u = 0; u < U; u++) {
                const int i = i0 + u * G::PARTS;
                i32 v = xa[u] * win[i] - xb[u] * win[2 * m - 1 - i];
                y[mt_ridx(m + i)] = v; range_add(rg, v);
                v = xc[u] * win[2 * m + i];
                if (i >= z) v += xd[u] * win[4 * m - 1 - i];
                y[mt_ridx(m - 1 - i)] = v; range_add(rg, v);
            }
        }
    } else {
        for (int i = part; i < N; i += G::PARTS) y[mt_ridx(i)] = 0;
    }
    mx[part * 32 + lane] = rg.mx; mn[part * 32 + lane] = rg.mn;
}

// ---- fold, 16-bit input, rows pulled from the caller's PCM: WARP = frame.  The 2N samples the fold of a frame needs
// (previous frame | current frame) are contiguous in the caller's [stream][sample] buffer, so they are brought into the
// frame's own row of X by ONE bulk asynchronous copy per row (cp.async.bulk + mbarrier, issued by mt_stage_rows in the
// kernel; 16-byte aligned rows only, others and the chunk's first frame -- whose previous frame is the staged history -- are
// filled in by the warp).  The fold then runs in place: every lane gathers the samples of its outputs into registers, the
// warp synchronises, the outputs overwrite the row.  This replaces the transposed copy of the whole PCM (k_stage) and the
// fold's strided global loads.
template <int N> struct MtRowIn { // where row r's staged samples sit: in[N + s] = sample s relative to the frame's start, s in [-N, N)
    LC3_HDM static i16 *at(i32 *X, int r) { return reinterpret_cast<i16 *>(X + r * MdctTile<N>::S) + ((r & 1) ? 4 : 0); } // 16-byte aligned
};

template <int N> LC3_HD void mt_fold_row(const LC3K_Batch &b, int r, bool valid, i32 *X, i32 *mx, i32 *mn, const i16 *win /* &rom_win[b.W_off], in shared memory in the kernel */)
{
    typedef MdctTile<N> G;
    constexpr int m = N / 2, K = (m + 31) / 32;
    i32 *y = X + r * G::S;
    const i16 *in = MtRowIn<N>::at(X, r) + N;
    const int memLen = b.mdct_mem_len, z = 2 * N - b.W_len;
    LC3_PERLANE(i32, v1, K);
    LC3_PERLANE(i32, v2, K);
    LC3_LANES(l)
#pragma unroll
        for (int k = 0; k < K; k++) {
            const int i = l + 32 * k;
            i32 a = 0, c = 0;
            if (valid && i < m) { // same three expressions as mdct_frame() / mt_fold()
                a = in[i - memLen] * win[i] - in[2 * m - 1 - i - memLen] * win[2 * m - 1 - i];
                c = in[2 * m - memLen + i] * win[2 * m + i];
                if (i >= z) c += in[4 * m - memLen - 1 - i] * win[4 * m - 1 - i]; // the window is zero below z (la_zeroes)
            }
            LC3_PL(v1, k, l) = a; LC3_PL(v2, k, l) = c;
        }
    LC3_LANES_END
    LC3_LANES(l)
        Range rg; range_init(rg);
#pragma unroll
        for (int k = 0; k < K; k++) {
            const int i = l + 32 * k;
            if (i < m) {
                y[mt_ridx(m + i)] = LC3_PL(v1, k, l); y[mt_ridx(m - 1 - i)] = LC3_PL(v2, k, l);
                range_add(rg, LC3_PL(v1, k, l)); range_add(rg, LC3_PL(v2, k, l));
            }
        }
        // reduced over the warp below: the slots of row r in the per-part arrays all receive the row's range
        LC3_PL(v1, 0, l) = rg.mx; LC3_PL(v2, 0, l) = rg.mn;
    LC3_LANES_END
    {
#if defined(__CUDA_ARCH__)
        const i32 rmx = __reduce_max_sync(0xffffffffu, v1[0]), rmn = __reduce_min_sync(0xffffffffu, v2[0]);
        const int l = (int)(threadIdx.x & 31);
        if (l < G::PARTS) { mx[l * 32 + r] = rmx; mn[l * 32 + r] = rmn; }
        __syncwarp();
#else
        i32 rmx = v1[0][0], rmn = v2[0][0];
        for (int l = 1; l < 32; l++) { rmx = imax(rmx, v1[0][l]); rmn = imin(rmn, v2[0][l]); }
        for (int p = 0; p < G::PARTS; p++) { mx[p * 32 + r] = rmx; mn[p * 32 + r] = rmn; }
#endif
    }
}

// what the bulk copy did not bring into row r: the previous frame of the chunk's first frame (staged history, tile-interleaved
// prefix of pcm_t), and whole rows whose address in the caller's buffer is not 16-byte aligned.  Warp-cooperative.
template <int N> LC3_HD void mt_fill_row(const LC3K_Batch &b, int f, int tile, int r, bool bulk_done, i32 *X)
{
    const int li = tile * LC3K_LANES + r;
    if (li >= b.n_lanes) return;
    i16 *in = MtRowIn<N>::at(X, r);
    const i16 *cur = (const i16 *)b.pcm + b.pcm_off[li] + (size_t)(b.frame0 + f) * N; // sample 0 of the frame
    SV<const i16> hist = tile_view((const i16 *)b.pcm_t, (size_t)tile, (size_t)(b.n_frames + 1) * N, r);
    LC3_LANES(l)
        if (f == 0) for (int s = l; s < N; s += 32) in[s] = hist[s];
        else if (!bulk_done) for (int s = l; s < N; s += 32) in[s] = cur[s - N];
        if (!bulk_done) for (int s = l; s < N; s += 32) in[N + s] = cur[s];
    LC3_LANES_END
}

// true when row r of (tile, f) was brought in by the bulk copy (mt_stage_rows applies the same test)
template <int N> LC3_HD bool mt_row_is_bulk(const LC3K_Batch &b, int f, int tile, int r)
{
    const int li = tile * LC3K_LANES + r;
    if (li >= b.n_lanes) return false;
    const i16 *cur = (const i16 *)b.pcm + b.pcm_off[li] + (size_t)(b.frame0 + f) * N;
    return (reinterpret_cast<uintptr_t>(f == 0 ? cur : cur - N) & 15) == 0;
}

#if defined(__CUDACC__)
__device__ __forceinline__ u32 mt_smem_addr(const void *p) { return (u32)__cvta_generic_to_shared(p); }
// Thread t < 32 of the block issues the bulk copy of row t and returns whether it did; every thread then waits on `bar`.
template <int N> __device__ __forceinline__ bool mt_stage_rows(const LC3K_Batch &b, int f, int tile, int t, i32 *X, unsigned long long *bar)
{
    bool issued = false;
    if (t == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mt_smem_addr(bar)) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (t < 32) {
        const int li = tile * LC3K_LANES + t;
        const i16 *src = nullptr;
        u32 bytes = 0;
        i16 *dst = MtRowIn<N>::at(X, t);
        if (li < b.n_lanes) {
            const i16 *cur = (const i16 *)b.pcm + b.pcm_off[li] + (size_t)(b.frame0 + f) * N;
            if (f == 0) { src = cur; bytes = 2 * N; dst += N; }      // the previous frame comes from the staged history
            else { src = cur - N; bytes = 4 * N; }
            if ((reinterpret_cast<unsigned long long>(src) & 15) != 0) bytes = 0; // mt_fill_row loads such rows
        }
        const u32 total = __reduce_add_sync(0xffffffffu, bytes);
        if (t == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mt_smem_addr(bar)), "r"(total) : "memory");
        __syncwarp();
        if (bytes) {
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(mt_smem_addr(dst)), "l"(src), "r"(bytes), "r"(mt_smem_addr(bar)) : "memory");
            issued = true;
        }
    }
    return issued;
}
__device__ __forceinline__ void mt_wait_rows(unsigned long long *bar)
{
    u32 ok;
    do {
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }"
                     : "=r"(ok) : "r"(mt_smem_addr(bar)) : "memory");
    } while (!ok);
}
#endif

// ---- block shift + pre-twiddle of frame fr: lane l takes pairs l, l+32, ...
template <int N> LC3_HD void mt_pre(i32 *X, const i32 *mx, const i32 *mn, int fr, int l, const MtTabs &T)
{
    typedef MdctTile<N> G;
    i32 *y = X + fr * G::S;
    const i32 s = mt_shift<N>(mx, mn, fr);
    const i16 *tw = T.tw;
    for (int kk = l; kk < G::M / 2; kk += 32) {
        const int lo = mt_ridx(2 * kk), hi = mt_ridx(N - 2 - 2 * kk);
        const cpx p = mt_ld2(y + lo), q = mt_ld2(y + hi);
        const i32 p0 = (i32)((u32)p.re << s), p1 = (i32)((u32)p.im << s), q0 = (i32)((u32)q.re << s), q1 = (i32)((u32)q.im << s);
        i32 a1, a2, a3, a4;
        cmul_half(a1, a2, q1, p0, tw[4 * kk], tw[4 * kk + 1]);
        cmul_half(a3, a4, q0, p1, tw[4 * kk + 2], tw[4 * kk + 3]);
        cpx o0 = {a2, a1}, o1 = {a4, -a3}; // cmul_half outputs are below 2^31 in magnitude
        mt_st2(y + lo, o0); mt_st2(y + hi, o1);
    }
}

// ---- leaf FFT i (0..7) of frame fr, in place: point m of the leaf is complex point i + 8m of the frame
template <int N> LC3_HD void mt_leaf(i32 *X, int fr, int i)
{
    typedef MdctTile<N> G;
    i32 *y = X + fr * G::S + 2 * i;
    fft_leaf<G::LEAF, G::SAT>(
        [&](int m) { cpx c = mt_ld2(y + 18 * m); c.re >>= 5; c.im >>= 5; return c; },
        [&](int m, cpx c) { mt_st2(y + 18 * m, c); });
}

// ---- column i (0..LEAF-1) of frame fr: output i of the 8 leaves = complex points 8i..8i+7
template <int N> LC3_HD void mt_col_compute(const i32 *X, int fr, int i, cpx *o, const MtTabs &T)
{
    typedef MdctTile<N> G;
    const i32 *y = X + fr * G::S + 18 * i;
    const i16 *W = T.W;
    cpx v[8];
#pragma unroll
    for (int k = 0; k < 8; k++) {
        const cpx a = mt_ld2(y + 2 * k);
        if (k == 0 || i == 0) {
            v[k].re = a.re >> 1;
            v[k].im = a.im >> 1;
        } else {
            const i32 c = W[G::SC * i + G::SC * k * G::LEAF - G::WOFF], d = W[G::SC * i + G::SC * k * G::LEAF + 1 - G::WOFF];
            // same column network as fft_n2 (lc3_fft.cuh): plain arithmetic by the bound proved there
            v[k].re = fs<G::SAT>(mul16<G::SAT>(a.re, c), mul16<G::SAT>(a.im, d)) >> 1;
            v[k].im = fa<G::SAT>(mul16<G::SAT>(a.re, d), mul16<G::SAT>(a.im, c)) >> 1;
        }
    }
    dft8<false, G::SAT>(v, o);
}
template <int N> LC3_HD void mt_col_store(i32 *X, int fr, int i, const cpx *o)
{
    typedef MdctTile<N> G;
    i32 *y = X + fr * G::S;
#pragma unroll
    for (int m = 0; m < 8; m++) mt_st2(y + mt_cidx(i + m * G::LEAF), o[m]);
}

// ---- post-twiddle of frame fr: lane l takes pairs l, l+32, ... (each pair touches only its own four words)
template <int N> LC3_HD void mt_post(i32 *X, int fr, int l, const MtTabs &T)
{
    typedef MdctTile<N> G;
    i32 *y = X + fr * G::S;
    const i16 *sn = T.sn;
    constexpr int step = (N % 3 == 0) ? 960 / N : 640 / N;
    constexpr int M2 = G::M / 2 - 1;
    for (int p = l; p <= M2; p += 32) {
        const int lo = mt_ridx(2 * p), hi = mt_ridx(N - 2 - 2 * p);
        const cpx L = mt_ld2(y + lo), H = mt_ld2(y + hi);
        i32 a1, a2, a3, a4;
        cpx oL, oH;
        if (p == M2) {
            cmul_half(a3, a4, L.im, L.re, sn[2 * p * step], sn[2 * p * step + 1]);
            a1 = mpy32x16k(H.re, 23170) >> 1;
            a2 = mpy32x16k(H.im, 23170) >> 1;
            oH.re = a1 + a2; oL.im = a1 - a2; oH.im = -a3; oL.re = a4; // |a1|, |a2| < 2^30; a3 != -2^31 (see cmul_half)
        } else {
            cmul_half(a3, a4, H.re, H.im, sn[2 * (p + 1) * step], sn[2 * (p + 1) * step + 1]);
            if (p == 0) { oH.im = -(L.im >> 1); oL.re = L.re >> 1; }
            else {
                cmul_half(a1, a2, L.im, L.re, sn[2 * p * step], sn[2 * p * step + 1]);
                oH.im = -a1; oL.re = a2;
            }
            oL.im = a3; oH.re = a4;
        }
        mt_st2(y + lo, oL); mt_st2(y + hi, oH);
    }
}

// ---- band energies (fx/per_band_energy_fx.c) while the spectrum is still in shared memory: thread t = (frame of
// the tile, part), bands part, part+8, ...; mantissas and exponents stay in shared memory until the store phase
// aligns them to the frame's largest exponent
struct MtRow { const i32 *y; LC3_HDM i32 operator[](int e) const { return y[mt_ridx(e)]; } };
template <int N> LC3_HD void mt_bands(const LC3K_Batch &b, int t, const i32 *X, i32 *ben, i32 *bexp, i32 *smx)
{
    typedef MdctTile<N> G;
    const int lane = t & 31, part = t >> 5;
    const MtRow row = {X + lane * G::S};
    const BandGeom g = band_geom(b.fs_idx, b.dms_idx);
    const i16 *bo = &rom_bands_offset[b.bands_off];
    i32 smax = -31;
    for (int band = part; band < b.nbands; band += G::PARTS) {
        i32 en;
        const i32 e = band_one(en, row, band, bo, g);
        ben[lane * G::BROW + band] = en; bexp[lane * G::BROW + band] = e;
        smax = imax(smax, e);
    }
    smx[part * 32 + lane] = smax;
}

// ---- store: thread t = (tile lane, part): spectrum, block exponent, aligned band energies and their exponent
template <int N, int BITS> LC3_HD void mt_store(const LC3K_Batch &b, int f, int tile, int t, const i32 *X, const i32 *mx, const i32 *mn,
                                      const i32 *ben, const i32 *bexp, const i32 *smx)
{
    typedef MdctTile<N> G;
    const int lane = t & 31, part = t >> 5, li = tile * LC3K_LANES + lane;
    if (li >= b.n_lanes) return;
    const size_t ft = frame_tile(b, f, tile);
    const i32 *y = X + lane * G::S;
    SV<i32> d = tile_view(b.spec, ft, N, lane);
    for (int e = part; e < b.yLen; e += G::PARTS) d[e] = y[mt_ridx(e)]; // nothing downstream reads a bin past yLen (400 >> dms_idx)
    const i32 x_exp = BITS == 16 ? 15 : tile_view((const i32 *)b.pout, ft, PO_N, lane)[PO_X_EXP];
    const i32 d_exp = s_add(s_add(s_sub(s_sub(x_exp, 2), mt_shift<N>(mx, mn, lane)), G::FFT_SCALE), 2);
    i32 smax = -31;
    for (int p = 0; p < G::PARTS; p++) smax = imax(smax, smx[p * 32 + lane]);
    SV<i32> eo = tile_view(b.ben, ft, BEN_N, lane);
    for (int band = part; band < b.nbands; band += G::PARTS)
        eo[band] = ben[lane * G::BROW + band] >> imin(s_sub(smax, bexp[lane * G::BROW + band]), 31);
    if (part == 0) {
        tile_view(b.fout, ft, FO_N, lane)[FO_D_EXP] = d_exp;
        eo[BEN_EXP] = s_add(s_shl(d_exp, 1), smax);
    }
}

} // namespace lc3k
