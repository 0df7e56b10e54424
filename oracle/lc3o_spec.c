/* ORACLE (test infrastructure, not product code).
 * Attack detector (fx/attack_detector_fx.c:13-119), per-band energy
 * (fx/per_band_energy_fx.c:17-228) and bandwidth detector (fx/detect_cutoff_warped_fx.c:15-118). */
#include "lc3o_internal.h"

/* fx/attack_detector_fx.c:13-119.  10 ms frames at 32/48 kHz only (4 blocks of 40 @16 kHz). */
void o_attack_detector(o_attdec_state *st, int attack_handling, const i16 *in, int N, int fs, i16 in_scaling)
{
    if (!attack_handling) return;
    i16 buf[2 + 160], *d = buf + 2;
    const int n16 = 160, nblocks = 4;
    i32 be[4];

    i16 s0 = s_add(o_headroom16(in, N, 0), in_scaling);
    i16 s1 = s_add(o_headroom16(st->filter_mem, 2, 16), st->scaling);
    i16 s2 = s_shr(s_add(s_add(o_headroom32(&st->acc_energy, 1, 32), s_shl(st->scaling, 1)), 1), 1);
    i16 ns = s_sub(s_min(s0, s_min(s1, s2)), 2);
    i16 rs = s_sub(ns, st->scaling);
    if (rs) {
        st->filter_mem[0] = s_shl(st->filter_mem[0], rs);
        st->filter_mem[1] = s_shl(st->filter_mem[1], rs);
        st->acc_energy = l_shl(st->acc_energy, s_shl(rs, 1));
    }
    st->scaling = ns;

    if (fs == 32000) {
        i16 ds = s_sub(1, s_sub(ns, in_scaling));
        for (int i = 0; i < n16; i++) d[i] = s_add(s_shr(in[2 * i], ds), s_shr(in[2 * i + 1], ds));
    } else { /* 48000 */
        i16 ds = s_sub(2, s_sub(ns, in_scaling));
        for (int i = 0; i < n16; i++)
            d[i] = s_add(s_shr(in[3 * i], ds), s_add(s_shr(in[3 * i + 1], ds), s_shr(in[3 * i + 2], ds)));
    }
    d[-2] = st->filter_mem[0];
    d[-1] = st->filter_mem[1];
    st->filter_mem[0] = d[n16 - 2];
    st->filter_mem[1] = d[n16 - 1];

    for (int i = n16 - 1; i >= 0; i--) { /* 3-tap high-pass, in place from the end */
        i32 t = l_mult(d[i], 12288);
        t = l_msu(t, d[i - 1], 16384);
        t = l_mac(t, d[i - 2], 4096);
        d[i] = hi16(t);
    }
    for (int j = 0; j < nblocks; j++) {
        be[j] = 0;
        for (int i = 0; i < 40; i++) be[j] = l_mac(be[j], d[i + j * 40], d[i + j * 40]);
    }

    st->detected = st->position >= 2;
    i16 pos = -1;
    for (int i = 0; i < nblocks; i++) {
        i16 l = (i16)(lo16(l_shr(be[i], 1)) & 0x7fff), h = hi16(be[i]); /* be/8.5 */
        i32 t = l_shr(l_mult0(l, 30840), 15);
        t = l_shr(l_mac0(t, h, 30840), 2);
        if (t > st->acc_energy) { pos = (i16)i; st->detected = 1; }
        st->acc_energy = l_max(l_shr(st->acc_energy, 2), be[i]);
    }
    st->position = pos;
}

/* |d|^2 of one bin after shifting left by s2 (<=15), as 16x16 product of the high word */
static inline i32 sq_hi(i32 v, i16 sh)
{
    i16 t = hi16(l_shl(v, sh));
    return l_mult0(t, t);
}
static inline i16 hr15(i32 v, i16 cur)
{
    return v != 0 ? s_min(cur, l_norm(v)) : cur;
}

/* fx/per_band_energy_fx.c:17-228 (linear == 0 variant only; the encoder never passes 1) */
void o_band_energy(i32 *en, i16 *en_exp, const i32 *d, i16 d_exp, const i16 *bo, int fs_idx, int nbands, int dms_idx)
{
    i16 be[LC3O_MAX_BANDS];
    int maxbin = 400 >> dms_idx;
    int one = rom_bands_one_max[dms_idx * 5 + fs_idx], two = rom_bands_two_max[dms_idx * 5 + fs_idx];
    int band = 0, i;

    for (; band < one; band++) { /* width-1 bands */
        i16 s2 = hr15(d[band], 15);
        en[band] = sq_hi(d[band], s2);
        be[band] = s_sub(1, s_shl(s2, 1));
    }
    i = one;
    for (; band < two; band++, i += 2) { /* width-2 bands */
        if (i + 1 >= maxbin) {
            if (i >= maxbin) {
                en[band] = 0;
                be[band] = 1 - 30;
            } else { /* reference reads d[band] here, not d[i] (per_band_energy_fx.c:117-128) */
                i16 s2 = hr15(d[band], 15);
                en[band] = sq_hi(d[band], s2);
                be[band] = s_sub(1, s_shl(s2, 1));
            }
        } else {
            i16 s2 = hr15(d[i + 1], hr15(d[i], 15));
            i32 e = l_min(sq_hi(d[i], s2), 0x3FFFFFFF);
            i16 t = hi16(l_shl(d[i + 1], s2));
            en[band] = l_shr(l_mac0(e, t, t), 1);
            be[band] = s_sub(1, s_shl(s2, 1));
        }
    }
    for (; band < nbands; band++) { /* wider bands */
        int stop = s_min(bo[band + 1], (i16)maxbin);
        i16 s1 = 15;
        for (int k = i; k < stop; k++) s1 = hr15(d[k], s1);
        int w = bo[band + 1] - bo[band];
        w = w < 0 ? 0 : (w > 31 ? 31 : w);
        i16 s2 = s_sub(s1, rom_bands_nrg_scale[w]);
        i32 e = 0;
        for (; i < stop; i++) {
            i16 t = hi16(l_shl(d[i], s2));
            e = l_mac0(e, t, t);
        }
        i = bo[band + 1];
        e = mpy32x16(e, rom_inv_int[w]);
        i16 s = l_norm(e);
        en[band] = l_shl(e, s);
        be[band] = s_sub(1, s_add(s_shl(s2, 1), s));
    }

    i16 smax = -31;
    for (band = 0; band < nbands; band++) smax = s_max(smax, be[band]);
    for (band = 0; band < nbands; band++) en[band] = l_shr(en[band], s_min(s_sub(smax, be[band]), 31));
    *en_exp = s_add(s_shl(d_exp, 1), smax);
}

/* fx/detect_cutoff_warped_fx.c:15-118 (called only for fs_idx > 0) */
i16 o_bw_detect(const i32 *en, i16 en_exp, int fs_idx, int dms_idx)
{
    const i16 *ws = &rom_bw_warp_start[(dms_idx * 4 + fs_idx - 1) * 4];
    const i16 *we = &rom_bw_warp_stop[(dms_idx * 4 + fs_idx - 1) * 4];
    const i16 *dist = &rom_bw_brickwall_dist[dms_idx * 4];
    int c = fs_idx;
    i16 below;
    do {
        c--;
        if (c < 0) break;
        int b = ws[c];
        i32 sm = en[b];
        i16 se = en_exp;
        for (b++; b <= we[c]; b++) sm = o_add_me(en[b], en_exp, sm, se, &se);
        i32 mean = mpy32x16(sm, rom_inv_int[we[c] - ws[c] + 1]);
        below = o_cmp_me(rom_bw_thresh_quiet[c], rom_bw_thresh_quiet_exp[0], mean, se);
    } while (below > 0);

    i16 bw = (i16)(c + 1);
    if (fs_idx - bw > 0) { /* brick-wall check just above the detected edge */
        int brick = 0, stop = ws[c + 1] + 1, dd = dist[c + 1];
        for (int b = stop; b >= stop - dd; b--) {
            i32 de = l_sub(mpy32x16(en[b - dd], rom_bw_thresh_brickwall[c + 1]), en[b]);
            if (de > 0) brick = 1;
            if (brick) break;
        }
        if (!brick) bw = (i16)fs_idx;
    }
    return bw;
}
