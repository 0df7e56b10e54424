// Per-band energy (fx/per_band_energy_fx.c:17-228) and bandwidth detector
// (fx/detect_cutoff_warped_fx.c:15-118), one frame per lane.
#pragma once
#include "lc3_math.cuh"

namespace lc3k {

// sh never exceeds the headroom of v at the call sites (it is a minimum of norm_l values), so L_shl cannot saturate
LC3_HD i32 sq_hi(i32 v, i32 sh) { i32 t = hi16(l_shl_ns(v, sh)); return t * t; }
LC3_HD i32 hr15(i32 v, i32 cur) { return v != 0 ? imin(cur, l_norm(v)) : cur; }

// Energy of ONE band before the bands are aligned: mantissa en, exponent returned (bands are independent of
// each other up to that alignment, which is what lets the tile kernel spread them over threads).
struct BandGeom { int maxbin, one, two; };
LC3_HD BandGeom band_geom(int fs_idx, int dms_idx)
{
    BandGeom g = {400 >> dms_idx, rom_bands_one_max[dms_idx * 5 + fs_idx], rom_bands_two_max[dms_idx * 5 + fs_idx]};
    return g;
}
template <class A> LC3_HD i32 band_one(i32 &en, A d, int band, const i16 *bo, const BandGeom &g)
{
    if (band < g.one) { // width-1 bands
        const i32 v = d[band];
        const i32 s2 = hr15(v, 15);
        en = sq_hi(v, s2);
        return s_sub(1, s_shl(s2, 1));
    }
    if (band < g.two) { // width-2 bands
        const int i = g.one + 2 * (band - g.one);
        if (i + 1 >= g.maxbin) {
            if (i >= g.maxbin) { en = 0; return 1 - 30; }
            // the reference reads d[band] here, not d[i] (per_band_energy_fx.c:117-128)
            const i32 v = d[band];
            const i32 s2 = hr15(v, 15);
            en = sq_hi(v, s2);
            return s_sub(1, s_shl(s2, 1));
        }
        const i32 v0 = d[i], v1 = d[i + 1];
        const i32 s2 = hr15(v1, hr15(v0, 15));
        const i32 a = imin(sq_hi(v0, s2), 0x3FFFFFFF);
        const i32 t = hi16(l_shl_ns(v1, s2));
        en = l_mac0(a, t, t) >> 1;
        return s_sub(1, s_shl(s2, 1));
    }
    // wider bands
    const int start = bo[band], stop = imin(bo[band + 1], g.maxbin);
    // minimum norm_l over the band's non-zero bins = norm_l of the OR of their magnitudes (v, or ~v when negative)
    u32 mag = 0;
    for (int k = start; k < stop; k++) { const i32 v = d[k]; mag |= (u32)(v ^ (v >> 31)); }
    const i32 s1 = hr15((i32)mag, 15);
    int w = bo[band + 1] - bo[band];
    w = w < 0 ? 0 : (w > 31 ? 31 : w);
    const i32 s2 = s_sub(s1, rom_bands_nrg_scale[w]);
    i64 a64 = 0; // non-negative terms: the saturating L_mac0 chain equals the clamped exact sum
    const int sl = s2 > 0 ? s2 : 0, sr = 16 + (s2 < 0 ? imin(-s2, 15) : 0); // the shift by s2 split once for the band, and the high half
    for (int i = start; i < stop; i++) {
        const i32 t = (i32)((u32)d[i] << sl) >> sr;
        a64 += (i64)t * t;
    }
    i32 a = a64 > LC3_I32_MAX ? LC3_I32_MAX : (i32)a64;
    a = mpy32x16(a, rom_inv_int[w]);
    const i32 s = l_norm(a);
    en = l_shl(a, s);
    return s_sub(1, s_add(s_shl(s2, 1), s));
}

// en[0..nbands) mantissas on the common exponent returned.  bo: bands_offset table of this config.
template <class A> LC3_HD i32 band_energy(i32 *en, A d, i32 d_exp, const i16 *bo, int fs_idx, int nbands, int dms_idx)
{
    i16 be[64];
    const BandGeom g = band_geom(fs_idx, dms_idx);
    i32 smax = -31;
    for (int band = 0; band < nbands; band++) {
        const i32 e = band_one(en[band], d, band, bo, g);
        be[band] = (i16)e; smax = imax(smax, e);
    }
    for (int band = 0; band < nbands; band++) en[band] = en[band] >> imin(s_sub(smax, be[band]), 31);
    return s_add(s_shl(d_exp, 1), smax);
}

// fx/detect_cutoff_warped_fx.c:15-118 (called only for fs_idx > 0)
LC3_HD i32 bw_detect(const i32 *en, i32 en_exp, int fs_idx, int dms_idx)
{
    const i16 *ws = &rom_bw_warp_start[(dms_idx * 4 + fs_idx - 1) * 4];
    const i16 *we = &rom_bw_warp_stop[(dms_idx * 4 + fs_idx - 1) * 4];
    const i16 *dist = &rom_bw_brickwall_dist[dms_idx * 4];
    int c = fs_idx;
    i32 below;
    do {
        c--;
        if (c < 0) break;
        int b = ws[c];
        i32 sm = en[b];
        i32 se = en_exp;
        for (b++; b <= we[c]; b++) sm = m_add_me(en[b], en_exp, sm, se, &se);
        i32 mean = mpy32x16(sm, rom_inv_int[we[c] - ws[c] + 1]);
        below = m_cmp_me(rom_bw_thresh_quiet[c], rom_bw_thresh_quiet_exp[0], mean, se);
    } while (below > 0);

    i32 bw = c + 1;
    if (fs_idx - bw > 0) { // brick-wall check just above the detected edge
        int brick = 0, stop = ws[c + 1] + 1, dd = dist[c + 1];
        for (int b = stop; b >= stop - dd; b--) {
            i32 de = l_sub(mpy32x16(en[b - dd], rom_bw_thresh_brickwall[c + 1]), en[b]);
            if (de > 0) { brick = 1; break; }
        }
        if (!brick) bw = fs_idx;
    }
    return bw;
}

} // namespace lc3k
