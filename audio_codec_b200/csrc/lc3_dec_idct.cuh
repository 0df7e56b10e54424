// DCT-IV of the decoder's inverse MDCT (fx/imdct_fx.c:72-103; fx/dct4_fx.c:16-138), frame-parallel.  The arithmetic is the
// encoder's DCT-IV (lc3_mdct.cuh / lc3_mdct_tile.cuh): the reference uses one dct_IV for both directions.
//   N >= 160   block per tile of 32 channel-frames, spectrum resident in shared memory (the phases of lc3_mdct_tile.cuh)
//   N <  160   lane per channel-frame, in place in the tile-interleaved spectrum
// Both leave the result ROW-MAJOR in b.yrow (one contiguous row per channel-frame) for the per-stream synthesis pass, with
// the block exponent and the up-shift of the windowing loops in DO_Y_E / DO_Y_S.
#pragma once
#include "lc3_dec_common.cuh"
#include "lc3_mdct_tile.cuh"

namespace lc3k {

// exponent bookkeeping around dct_IV (fx/imdct_fx.c:76-103).  y_s0: getScaleFactor32_0 of the spectrum; head: getScaleFactor32
// of the DCT output.
LC3_HD void idct_exponents(int N, i32 q_exp, i32 y_s0, i32 fft_scale, i32 head, i32 &y_e, i32 &y_s)
{
    if (y_s0 < 32) {
        y_e = s_sub(q_exp, y_s0);
        y_e = s_add(s_add(y_e, fft_scale), 2);
        y_s = s_sub(head, 1);
        y_e = s_sub(y_e, y_s + 3);
        if (N <= 20) y_e = s_add(y_e, 2);
        else if (N <= 120) y_e = s_add(y_e, 1);
    } else {
        y_e = 0; y_s = y_s0;
    }
}

template <int N> LC3_HD void idct_lane_body(const LC3D_Batch &b, int f, int tile, int lane)
{
    const int li = tile * LC3K_LANES + lane;
    if (li >= b.n_lanes) return;
    const size_t ft = dframe_tile(b, f, tile);
    SV<i32> d = tile_view(b.spec, ft, N, lane), w = tile_view(b.work, ft, N, lane);
    SV<i32> dout = tile_view(b.dout, ft, DO_N, lane);
    const i32 y_s0 = dout[DO_Y_S];
    i32 fft_scale = 0, head = 0, y_e, y_s;
    if (y_s0 < 32) {
        fft_scale = dct4_frame<N>(d, w, y_s0);
        head = m_headroom32(d, N, 0);
    }
    idct_exponents(N, dout[DO_Q_EXP], y_s0, fft_scale, head, y_e, y_s);
    dout[DO_Y_E] = y_e; dout[DO_Y_S] = y_s;
}

// ---- tile form: load (thread = (tile lane, part)), the DCT phases of lc3_mdct_tile.cuh, store (warp = channel-frame)
template <int N> LC3_HD void mt_dec_load(const LC3D_Batch &b, int f, int tile, int t, i32 *X, i32 *mx, i32 *mn)
{
    typedef MdctTile<N> G;
    const int lane = t & 31, part = t >> 5, li = tile * LC3K_LANES + lane;
    i32 *y = X + lane * G::S;
    Range rg; range_init(rg);
    if (li < b.n_lanes) {
        SV<const i32> d = tile_view((const i32 *)b.spec, dframe_tile(b, f, tile), N, lane);
        for (int e = part; e < N; e += G::PARTS) { const i32 v = d[e]; y[mt_ridx(e)] = v; range_add(rg, v); }
    } else {
        for (int e = part; e < N; e += G::PARTS) y[mt_ridx(e)] = 0;
    }
    mx[part * 32 + lane] = rg.mx; mn[part * 32 + lane] = rg.mn;
}
// red: 64 words of warp-private scratch
template <int N> LC3_HD void mt_dec_store(const LC3D_Batch &b, int f, int tile, int fr, const i32 *X, i32 *red)
{
    typedef MdctTile<N> G;
    const int li = tile * LC3K_LANES + fr;
    if (li >= b.n_lanes) return;
    const i32 *y = X + fr * G::S;
    i32 *row = b.yrow + dframe_row(b, f, li) * N;
    LC3_LANES(l)
        Range rg; range_init(rg);
        for (int e = l; e < N; e += 32) { const i32 v = y[mt_ridx(e)]; row[e] = v; range_add(rg, v); }
        red[l] = rg.mx; red[32 + l] = rg.mn;
    LC3_LANES_END
    Range rr;
    rr.mx = warp_max(red); rr.mn = warp_min(red + 32);
    LC3_SYNC;
    LC3_LANE0
        SV<i32> dout = tile_view(b.dout, dframe_tile(b, f, tile), DO_N, fr);
        i32 y_e, y_s;
        idct_exponents(N, dout[DO_Q_EXP], dout[DO_Y_S], G::FFT_SCALE, range_headroom32(rr, 0), y_e, y_s);
        dout[DO_Y_E] = y_e; dout[DO_Y_S] = y_s;
    LC3_LANE0_END
}

} // namespace lc3k
