// Rate-loop pass, ONE WARP PER CHANNEL-STREAM (fx/enc_lc3.c:180-192): the second half of
// processEstimateGlobalGain_fx (fx/estimate_global_gain_fx.c:44-59, :112-204) and the first quantisation
// pass' bit count (fx/quantize_spec_fx.c:15-221), which feed targetBitsOff / mem_targetBits / mem_specBits
// from one frame of a stream to the next.  Lanes split the quadruples, the bins and the 2-tuples.
//
// The context of a 2-tuple in the reference, c = ((c & 0xf) << 4) + t(previous tuple), is a pure function of
// the two previous tuples' magnitudes class t (< 16), so all contexts and table look-ups of a frame are
// computed in parallel and the bit count is one warp reduction (mode -1 needs no prefix sums).
#pragma once
#include "lc3_stages.cuh"
#include "lc3_warp.cuh"
#include "lc3_rate.cuh"

namespace lc3k {

constexpr int RATE_MAXT = (LC3K_MAX_N / 2 + 31) / 32; // tuples per lane (8; yLen <= 400 uses 7)

struct RateShared {
    WarpRed red;
    u8 tv[LC3K_MAX_N / 2 + 2];   // magnitude class of each tuple, offset by 2 (tv[0] = tv[1] = 0)
};

// sync_threads > 0 (kernel only): the block's active warps meet at a barrier before every frame (see pitch_warp_body)
LC3_HD void rate_warp_body(const LC3K_Batch &b, int li, RateShared &sh, const u8 *ari_lut, const uint16_t *ari_bitstab,
                           int sync_threads = 0)
{
    if (li >= b.n_lanes) return;
    const int tile = li >> 5, tl = li & 31;
    const int slot = b.slot[li];
    SV<i32> st = slot_view(b.st32, slot, ST32_N);
    SV<const i32> cfg = slot_view(b.cfg, slot, CFG_N);
    const i32 tbits_init = cfg[CFG_TBITS_INIT], total_bits = cfg[CFG_TOTAL_BITS], qgain_off = cfg[CFG_QGAIN_OFF];
    i32 tbits_off = st[ST_TBITS_OFF], mem_tbits = st[ST_MEM_TBITS], mem_specbits = st[ST_MEM_SPECBITS];
    const int nt = b.yLen, lg4 = nt >> 2, nt_half = nt >> 1;
    const i32 rateFlag = s_sub(total_bits, s_add(160, s_imult(b.fs_idx, 160))) > 0 ? 2 << 8 : 0;

    for (int f = 0; f < b.n_frames; f++) {
#if defined(__CUDA_ARCH__)
        if (sync_threads > 0) asm volatile("bar.sync 1, %0;" ::"r"(sync_threads) : "memory");
#endif
        const size_t ft = frame_tile(b, f, tile);
        SV<const i32> fo = tile_view((const i32 *)b.fout, ft, FO_N, tl);
        SV<const i32> po = tile_view((const i32 *)b.pout, ft, PO_N, tl);
        SV<const i32> d = tile_view((const i32 *)b.spec, ft, b.N, tl);
        SV<const i32> en = tile_view((const i32 *)b.en4, ft, b.N / 4, tl);
        const i32 x_e = fo[FO_D_EXP];
        i32 x_max = fo[FO_X_MAX];
        const i32 tbits_quant = s_sub(tbits_init, s_add(fo[FO_TNS_BITS], po[PO_LTPF_BITS]));

        // ---- rate-loop feedback (estimate_global_gain_fx.c:44-59)
        if (mem_tbits < 0) {
            tbits_off = 0;
        } else {
            i32 t = l_add(tbits_off, l_dep_h(s_sub(mem_tbits, mem_specbits)));
            t = l_min(40 << 16, l_max(-(40 << 16), t));
            tbits_off = l_add(mpy32x16k(tbits_off, 26214), mpy32x16k(t, 6554));
        }
        mem_tbits = tbits_quant;
        const i32 nbitsSQ = s_add(tbits_quant, l_round(tbits_off));

        i32 qgain, qgain_min;
        if (x_max == 0) {
            qgain_min = qgain_off;
            qgain = 0;
            mem_tbits = -1;
        } else {
            x_max = l_add(m_log2(x_max), l_shl_p(x_e, 25));
            x_max = l_sub(mpy32x32k(x_max, 0x436E439A), 0xFCDD6B4);
            qgain_min = lo16(l_add(x_max, (1 << 21) + (1 << 11)) >> 21);
            qgain_min = s_max(qgain_off, s_min(s_add(255, qgain_off), qgain_min));

            // ---- bisection on the gain index (estimate_global_gain_fx.c:127-188): lane = quadruple.
            // The reference walks the quadruples from the top with an "all zero so far" flag: a quadruple
            // below the threshold costs 0x3C7AE only if a higher one is above it.  All terms are >= 0, so
            // the saturating running sum equals the clamped total.
            i32 offset = l_dep_h(s_add(255, qgain_off));
            const i32 target = l_shl_p(l_mult(0x7D71, nbitsSQ), 1);
            i32 fac = 0x1000000;
            // |en| < 2^25 by construction (gain_energies: (mpy32x16k(t, 0x436E) >> 6) + 0x9CCCD for any 32-bit t) and
            // |offset| < 2^24 (a gain index in [0, 255] << 16 minus at most 255 << 16): en - offset cannot saturate
            // and a quadruple's cost is below 2^27 (tests/hostsim/primcheck.cpp: rate_bounds).
            // The four quadruples of a lane stay in registers over the eight steps.
            LC3_PERLANE(i32, enr, 4);
            LC3_LANES(l)
#pragma unroll
                for (int k = 0; k < 4; k++) LC3_PL(enr, k, l) = l + 32 * k < lg4 ? en[l + 32 * k] : -(1 << 30);
            LC3_LANES_END
            for (int it = 0; it < 8; it++) {
                fac = fac >> 1;
                offset = l_sub(offset, fac);
                // One pass: every quadruple at or above the threshold adds its own cost; the ones below it cost
                // 0x3C7AE each, but only underneath the highest one above it -- top + 1 - (number above) of those,
                // so each one above carries -0x3C7AE and the reduction needs no count.
                LC3_LANES(l)
                    i32 top = -1, est = 0; // four terms below 2^27
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        const i32 t = LC3_PL(enr, k, l) - offset;
                        if (t >= 0x9CCCD) {
                            top = l + 32 * k;
                            est += t - 0x3C7AE + imax(t - 0x460000, 0);
                        }
                    }
                    sh.red.a[l] = est & 0xFFFF; sh.red.b[l] = est >> 16; sh.red.d[l] = top;
                LC3_LANES_END
                const i32 top = warp_max(sh.red.d);
                i64 est64 = ((i64)warp_sum(sh.red.b) << 16) + warp_sum(sh.red.a);
                if (top >= 0) est64 += (i64)0x3C7AE * (top + 1);
                const i32 est = clamp64(est64);
                if (est > target && top >= 0) offset = l_add(offset, fac);
            }
            const i32 g = hi16(offset);
            if (g < qgain_min) mem_tbits = -1;
            qgain = s_sub(s_max(qgain_min, g), qgain_off);
        }
        i32 gain, gain_e;
        gain_from_index(s_add(qgain, qgain_off), gain, gain_e);

        // ---- first quantisation pass, bit count only (quantize_spec_fx.c:53-221, mode -1): lane = bin, then tuple
        i32 ge = gain_e;
        const i32 inv = m_inv16(gain, &ge);
        const i32 s = s_sub(s_add(x_e, ge), 15);
        // lane l owns tuples l, l + 32, ...: their magnitudes stay in its registers between the two phases
        LC3_PERLANE(i32, ab, RATE_MAXT);
        LC3_LANES(l)
            i32 last = 0;
#pragma unroll
            for (int m = 0; m < RATE_MAXT; m++) {
                const int k = 2 * l + 64 * m;
                if (k >= nt) break;
                const i32 a1 = s_abs((i16)quant_bin(d[k], inv, s)), b1 = s_abs((i16)quant_bin(d[k + 1], inv, s)), mx = imax(a1, b1);
                LC3_PL(ab, m, l) = a1 | (b1 << 16);
                // magnitude class of the tuple (the value the reference folds into the next two contexts)
                // (1 | a1 + b1 + 1 | 2 * (a1 / 2 + b1 / 2) + 1 | 12 + min(lev, 3) for mx = 0 | < 4 | < 8 | above, with lev = 0, 0, 1, >= 2)
                const i32 lev = imax(0, 13 - s_norm(mx | 1));
                const i32 t = lev < 2 ? (((a1 >> lev) + (b1 >> lev)) << lev) + 1 : 12 + imin(lev, 3);
                sh.tv[2 + (k >> 1)] = (u8)t;
                if (mx != 0 && k >= 2) last = k + 2; // find_last_nz_pair: the pair below bin 2 is not looked at
            }
            sh.red.a[l] = last;
            if (l < 2) sh.tv[l] = 0;
        LC3_LANES_END
        const i32 lastnz = imax(warp_max(sh.red.a), 2);
        LC3_LANES(l)
            i32 nb = 0;
#pragma unroll
            for (int m = 0; m < RATE_MAXT; m++) {
                const int k = 2 * l + 64 * m;
                if (k >= lastnz) break;
                const int j = k >> 1;
                i32 t = ((i32)sh.tv[j] << 4) + sh.tv[j + 1] + rateFlag; // tv[j] = class of tuple j-2, tv[j+1] = of j-1
                if (k > nt_half) t += 1 << 8;
                const i32 a1 = LC3_PL(ab, m, l) & 0xffff, b1 = LC3_PL(ab, m, l) >> 16, mx = imax(a1, b1);
                // One straight path for every magnitude class (the lanes of a warp hold all of them at once): `lev` escape
                // symbols (row t + 1024 * min(q, 3), symbol 16) and lev LSB pairs, the MSB symbol, the signs.
                // 13 - norm_s(mx) for mx >= 4, 0 below (an all-zero tuple codes symbol 0 of its own row).
                const i32 lev = imax(0, 13 - s_norm(mx | 1));
                nb += (lev << 12) + (((a1 != 0) + (b1 != 0)) << 11);
                if (lev > 0) nb += ari_bits(ari_lut, ari_bitstab, t, 16);
                if (lev > 1) nb += ari_bits(ari_lut, ari_bitstab, t + 1024, 16);
                if (lev > 2) nb += ari_bits(ari_lut, ari_bitstab, t + 2048, 16);
                if (lev > 3) nb += (lev - 3) * ari_bits(ari_lut, ari_bitstab, t + 3072, 16);
                nb += ari_bits(ari_lut, ari_bitstab, t + (imin(lev, 3) << 10), (a1 >> lev) + ((b1 >> lev) << 2));
            }
            sh.red.a[l] = nb;
        LC3_LANES_END
        const i32 nbq = warp_sum(sh.red.a); // Q11 bits; a frame holds < 2^15 bits, so no 32-bit overflow
        const i32 nBits = s_add(lo16((nbq - 1) >> 11), 1);
        mem_specbits = nBits;

        LC3_LANES(l)
            if (l == 0) {
                SV<i32> ro = tile_view(b.rout, ft, RO_N, tl);
                ro[RO_QGAIN] = qgain; ro[RO_QGAIN_MIN] = qgain_min; ro[RO_NBITS1] = nBits; ro[RO_TBITS_QUANT] = tbits_quant;
                ro[RO_LASTNZ] = lastnz; // of the first pass: the final one unless the gain is adjusted (lc3_quant_tile.cuh)
            }
        LC3_LANES_END
    }
    LC3_LANES(l)
        if (l == 0) { st[ST_TBITS_OFF] = tbits_off; st[ST_MEM_TBITS] = mem_tbits; st[ST_MEM_SPECBITS] = mem_specbits; }
    LC3_LANES_END
}

} // namespace lc3k
