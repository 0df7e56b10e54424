// Final quantisation of one TILE (32 channel-frames) per thread block (fx/enc_lc3.c:196-215):
// processAdjustGlobalGain_fx (fx/adjust_global_gain_fx.c:14-83), then processQuantizeSpec_fx in mode -1 / 0 / 1
// (fx/quantize_spec_fx.c:15-553): quantised bins, last non-zero pair, bit count, truncation point, LSB mode.
//
// It replaces the lane-per-frame quantize<true>() that used to open the back-end kernel, where the per-tuple
// escape loops of 32 different frames diverged (14 of 32 lanes active on average).  Mapping:
//   setup     thread = tile lane: gain adjustment and the frame's quantiser scalars
//   quantise  thread = (tile lane, 1/8 of the bins): coalesced read of the spectrum, bins to shared memory
//   analyse   warp = frame (four frames per warp): last non-zero pair, tuple classes, per-tuple bit costs with
//             every context known up front (a context is a function of the two previous tuples' class, see
//             lc3_rate_warp.cuh), block-wise prefix sums for the truncation point
//   store     thread = (tile lane, 1/8 of the bins): coalesced write of the (truncated) bins
#pragma once
#include "lc3_stages.cuh"
#include "lc3_warp.cuh"
#include "lc3_rate.cuh"

namespace lc3k {

constexpr int QT_THREADS = 256, QT_PARTS = QT_THREADS / 32;
constexpr int QT_ROW = LC3K_MAX_N + 2;        // halfwords per frame row; ROW/2 odd: tile-lane-major access is conflict-free
constexpr int QT_MAXT = (LC3K_MAX_N / 2 + 31) / 32; // tuples per lane in the block-wise pass

struct QuantFrame { // per frame of the tile
    i32 inv, s, mode, rateFlag, target, total_bits, qgain, valid, nbits1;
    i32 lastnz, lastnz_out, lsbMode, nBits2;
};
struct QuantRed { i32 a[32], b[32], c[32], d[32]; };  // reduction slots (lc3_warp.cuh)
struct QuantWarp { QuantRed red; i32 nb[LC3K_MAX_N / 2]; u8 tv[LC3K_MAX_N / 2 + 2]; }; // nb: bit cost of each tuple
struct QuantShared {
    i16 xq[32 * QT_ROW];
    QuantFrame fr[32];
    QuantWarp w[QT_PARTS];
};

// ---- setup: thread = tile lane
LC3_HD void qt_setup(const LC3K_Batch &b, int f, int tile, int lane, QuantFrame &q)
{
    const int li = tile * LC3K_LANES + lane;
    q.valid = li < b.n_lanes;
    if (!q.valid) { q.inv = 0; q.s = 0; q.mode = -1; q.rateFlag = 0; q.target = 0; q.total_bits = 0; q.qgain = 0; q.nbits1 = 0; return; }
    SV<const i32> cfg = slot_view(b.cfg, b.slot[li], CFG_N);
    const size_t ft = frame_tile(b, f, tile);
    SV<const i32> fo = tile_view((const i32 *)b.fout, ft, FO_N, lane);
    SV<const i32> ro = tile_view((const i32 *)b.rout, ft, RO_N, lane);
    const i32 total_bits = cfg[CFG_TOTAL_BITS], qgain_off = cfg[CFG_QGAIN_OFF], tbits_quant = ro[RO_TBITS_QUANT];
    i32 qgain = ro[RO_QGAIN], gain, gain_e;
    gain_from_index(s_add(qgain, qgain_off), gain, gain_e);
    const int change = adjust_gain(qgain, ro[RO_QGAIN_MIN], qgain_off, gain, gain_e, tbits_quant, ro[RO_NBITS1], b.fs_idx);
    i32 ge = gain_e;
    q.inv = m_inv16(gain, &ge);
    q.s = s_sub(s_add(fo[FO_D_EXP], ge), 15);
    i32 mode = change ? 0 : -1; // fx/enc_lc3.c:203-215
    if (mode == 0 && s_sub(total_bits, s_add(480, s_imult(b.fs_idx, 160))) >= 0) mode = 1;
    q.mode = mode;
    q.rateFlag = s_sub(total_bits, s_add(160, s_imult(b.fs_idx, 160))) > 0 ? 2 << 8 : 0;
    q.target = tbits_quant; q.total_bits = total_bits; q.qgain = qgain; q.nbits1 = ro[RO_NBITS1];
    q.lastnz = ro[RO_LASTNZ]; // the first pass' (lc3_rate_warp.cuh); replaced below when the gain changed
}

// ---- quantise: thread t = (tile lane, part)
LC3_HD void qt_quantise(const LC3K_Batch &b, int f, int tile, int t, QuantShared &sh)
{
    const int lane = t & 31, part = t >> 5;
    const QuantFrame &q = sh.fr[lane];
    i16 *xq = sh.xq + lane * QT_ROW;
    if (q.valid) {
        SV<const i32> d = tile_view((const i32 *)b.spec, frame_tile(b, f, tile), b.N, lane);
        for (int e = part; e < b.yLen; e += QT_PARTS) xq[e] = (i16)quant_bin(d[e], q.inv, q.s);
    } else {
        for (int e = part; e < b.yLen; e += QT_PARTS) xq[e] = 0;
    }
}

// bit cost (Q11) of tuple (a1, b1) in context t, and what it adds to the diverted LSB plane (mode 1)
LC3_HD i32 qt_tuple_bits(i32 t, i32 a1, i32 b1, i32 mode, const u8 *lut, const uint16_t *bitstab, i32 &lsb)
{
    // one straight path for every magnitude class, as in rate_warp_body: lev = 0 below 4, 1 below 8, 13 - norm_s(mx) above
    const i32 mx = imax(a1, b1);
    const i32 lev = imax(0, 13 - s_norm(mx | 1));
    const i32 dv = (mode > 0 && lev > 0) ? 1 : 0; // the lowest plane is diverted: its bits are counted apart, signs of +-1 go with it
    i32 nb = ((lev - dv) << 12) + ((((a1 >> dv) != 0) + ((b1 >> dv) != 0)) << 11);
    lsb = dv ? 2 + (a1 == 1) + (b1 == 1) : 0;
    if (lev > 0) nb += ari_bits(lut, bitstab, t, 16);
    if (lev > 1) nb += ari_bits(lut, bitstab, t + 1024, 16);
    if (lev > 2) nb += ari_bits(lut, bitstab, t + 2048, 16);
    if (lev > 3) nb += (lev - 3) * ari_bits(lut, bitstab, t + 3072, 16);
    nb += ari_bits(lut, bitstab, t + (imin(lev, 3) << 10), (a1 >> lev) + ((b1 >> lev) << 2));
    return nb;
}

// ---- analyse one frame: warp-cooperative.  All bit sums are far below 2^31 (a frame holds < 2^15 bits in Q11), so
// the reference's saturating L_add chain is plain addition here.
LC3_HD void qt_analyse(QuantFrame &q, i16 *xq, int nt, QuantWarp &w, const u8 *lut, const uint16_t *bitstab)
{
    const i32 mode = q.mode, rateFlag = q.rateFlag;
    const int nt_half = nt >> 1;
    // Without a gain change (mode -1) this pass repeats the first one (lc3_rate_warp.cuh) bin for bin: nothing is truncated,
    // the last non-zero pair and the bit count are the ones that pass left in RO_LASTNZ / RO_NBITS1.
    if (mode < 0) {
        LC3_LANES(l)
            if (l == 0) { q.lastnz_out = q.lastnz; q.nBits2 = q.nbits1; q.lsbMode = 0; }
        LC3_LANES_END
        return;
    }
    LC3_LANES(l)
        i32 last = 0;
        for (int i = l; i < nt; i += 32)
            if (xq[i] != 0 && i >= 2) last = (i | 1) + 1; // find_last_nz_pair (fx/ari_codec.c:1269-1308)
        w.red.a[l] = last;
        if (l < 2) w.tv[l] = 0;
    LC3_LANES_END
    const i32 lastnz = imax(warp_max(w.red.a), 2);
    LC3_LANES(l) // magnitude class of every tuple (what the reference folds into the next two contexts)
        for (int k = 2 * l; k < lastnz; k += 64) {
            const i32 a1 = s_abs(xq[k]), b1 = s_abs(xq[k + 1]), mx = imax(a1, b1);
            const i32 lev = imax(0, 13 - s_norm(mx | 1));
            const i32 t = lev < 2 ? (((a1 >> lev) + (b1 >> lev)) << lev) + 1 : 12 + imin(lev, 3);
            w.tv[2 + (k >> 1)] = (u8)t;
        }
    LC3_LANES_END
    i32 lastnz2 = lastnz, nBits2 = q.nbits1, tot = q.nbits1;
    {
        // block-wise: lane l owns tuples [l*T, (l+1)*T); their costs are kept for the prefix pass
        const int ntup = lastnz >> 1, T = (ntup + 31) >> 5;
        LC3_LANES(l)
            i32 sum = 0, lsb_tot = 0;
            for (int j = l * T; j < imin(ntup, (l + 1) * T); j++) {
                const int k = 2 * j;
                i32 t = ((i32)w.tv[j] << 4) + w.tv[j + 1] + rateFlag;
                if (k > nt_half) t += 1 << 8;
                i32 lsb;
                const i32 nb = qt_tuple_bits(t, s_abs(xq[k]), s_abs(xq[k + 1]), mode, lut, bitstab, lsb);
                w.nb[j] = nb;
                sum += nb; lsb_tot += lsb;
            }
            w.red.a[l] = sum; w.red.b[l] = lsb_tot;
        LC3_LANES_END
        const i32 nb_all = warp_sum(w.red.a), nbits_lsb = warp_sum(w.red.b);
        const i32 target32 = l_shl_p(q.target, 11);
        // truncation point: last non-zero tuple at which the running count is still within the target
        LC3_LANES(l)
            i32 run = -target32;
            for (int k = 0; k < l; k++) run += w.red.a[k];
            i32 best_k = 0, best_nb = 0;
            for (int j = l * T; j < imin(ntup, (l + 1) * T); j++) {
                run += w.nb[j];
                if (w.tv[2 + j] != 1 && run <= 0) { best_k = 2 * j + 2; best_nb = run; } // class 1 = all-zero tuple
            }
            w.red.c[l] = best_k; w.red.d[l] = best_nb;
        LC3_LANES_END
        const i32 bk = warp_max(w.red.c);
        i32 nb2 = 0;
        lastnz2 = 2;
        if (bk > 0) {
            lastnz2 = bk;
            LC3_LANES(l)
                w.red.a[l] = w.red.c[l] == bk ? w.red.d[l] : LC3_I32_MIN; // exactly one lane owns tuple bk/2 - 1
            LC3_LANES_END
            nb2 = warp_max(w.red.a);
        }
        tot = s_add(lo16(l_sub(nb_all, 1) >> 11), 1);
        nBits2 = s_add(lo16(l_sub(l_add(nb2, target32), 1) >> 11), 1);
        if (mode > 0) { tot = s_add(tot, nbits_lsb); nBits2 = s_add(nBits2, nbits_lsb); }
    }
    LC3_LANES(l)
        if (l == 0) {
            q.lastnz = lastnz; q.lastnz_out = lastnz2; q.nBits2 = nBits2;
            q.lsbMode = (mode > 0 && s_sub(tot, q.target) > 0) ? 1 : 0;
        }
    LC3_LANES_END
}

// ---- store: thread t = (tile lane, part)
LC3_HD void qt_store(const LC3K_Batch &b, int f, int tile, int t, const QuantShared &sh)
{
    const int lane = t & 31, part = t >> 5;
    const QuantFrame &q = sh.fr[lane];
    if (!q.valid) return;
    const size_t ft = frame_tile(b, f, tile);
    const i16 *xq = sh.xq + lane * QT_ROW;
    SV<i16> o = tile_view(b.xq, ft, b.N, lane);
    for (int e = part; e < b.yLen; e += QT_PARTS) o[e] = (e >= q.lastnz_out && e < q.lastnz) ? (i16)0 : xq[e];
    if (part == 0) {
        SV<i32> ro = tile_view(b.rout, ft, RO_N, lane);
        ro[RO_QGAIN] = q.qgain; ro[RO_LASTNZ] = q.lastnz_out; ro[RO_LSBMODE] = q.lsbMode; ro[RO_NBITS2] = q.nBits2;
    }
}

} // namespace lc3k
