// Time-domain half of the batched LC3plus decoder, ONE WARP PER CHANNEL-STREAM, frames in order: windowing and overlap-add
// of the inverse MDCT (fx/imdct_fx.c:105-166), LTPF synthesis (fx/ltpf_decoder_fx.c:22-366) and output scaling
// (fx/dec_lc3.c:289-311), for good frames (bfi = 0; no concealment in this library).
//
// This is the chain whose state (overlap memory and its exponent, LTPF input/output memories, their exponent, the previous
// frame's pitch and gain) makes frame n+1 wait for frame n.  The parallelism is inside the frame: the 32 lanes split the
// sample loops, with the stream's memories resident in the warp's shared-memory block for the whole chunk.  The DCT-IV
// output of every frame arrives ROW-MAJOR (one contiguous row per channel-frame, written by the frame-parallel DCT pass), so
// the warp's loads are coalesced; the PCM goes straight to the caller's buffer.  Phases and reductions follow lc3_warp.cuh.
#pragma once
#include "lc3_dec_common.cuh"

namespace lc3k {

#define LC3D_BY_LEN (LC3D_MAX_LTPF_Y + 3 + LC3K_MAX_N)
struct SynthShared {
    WarpRed red;
    i16 xf[2 * LC3K_MAX_N];                    // time signal of the frame: N samples (after the LTPF: the output), then the new overlap memory
    i16 ola[LC3K_MAX_N];                       // overlap-add memory
    i16 bx[LC3D_MAX_LTPF_X + 2 + LC3K_MAX_N];  // LTPF input: ltpf_mem_x | current frame
    i16 by[LC3D_BY_LEN];                       // LTPF output: ltpf_mem_y | current frame
    i16 z[128 + 16];                           // transition buffer of a pitch change (ltpf_decoder_fx.c:222-226)
};

// 16-bit headroom of a block spread over the warp (getScaleFactor16 / getScaleFactor16_0, fx/basop_util.c:286-357)
LC3_HD i32 warp_headroom16(const i16 *x, int n, int zero_val, WarpRed &red)
{
    LC3_LANES(l)
        Range r; range_init(r);
        for (int k = l; k < n; k += 32) range_add(r, x[k]);
        red.a[l] = r.mx; red.b[l] = r.mn;
    LC3_LANES_END
    Range rr;
    rr.mx = warp_max(red.a); rr.mn = warp_min(red.b);
    LC3_SYNC;
    return range_headroom16(rr, zero_val);
}
// Scale_sig (fx/basop_util.c:571-595) spread over the warp
LC3_HD void warp_scale_sig(i16 *x, int n, i32 e)
{
    if (e == 0) return;
    const i32 t = e < 0 ? s_shl(-32768, s_max(e, -15)) : 0;
    LC3_LANES(l)
        for (int k = l; k < n; k += 32) x[k] = (i16)scale_val(x[k], e, t);
    LC3_LANES_END
}

// x[k] = x[k + n] for k < len (the reference's basop_memmove towards the front), 32 elements at a time through registers: a
// chunk only reads elements no earlier chunk has written (n > 0), and inside a chunk every lane has read before any writes
LC3_HD void warp_shift_down(i16 *x, int len, int n)
{
    for (int k0 = 0; k0 < len; k0 += 32) {
        LC3_PERLANE(i16, t, 1);
        LC3_LANES(l)
            if (k0 + l < len) LC3_PL(t, 0, l) = x[k0 + l + n];
        LC3_LANES_END
        LC3_LANES(l)
            if (k0 + l < len) x[k0 + l] = LC3_PL(t, 0, l);
        LC3_LANES_END
    }
}

// ltpf_synth_filter (fx/ltpf_decoder_fx.c:268-366).  out[j] depends on out[j - pitch_int + shift - l], l >= 0: the shortest
// lag is pitch_int - shift >= 19 samples at 8 kHz and >= 39 from 16 kHz up (pitch >= 32 samples at 12.8 kHz,
// ltpf_decoder_fx.c:66-88), so 16 (8 kHz) or 32 consecutive outputs are independent of each other and are computed by the
// lanes side by side.  The fade factor alpha advances by a constant step from 0 or 0x7FFF and stays inside [0, 0x7FFF]
// over the 2.5 ms it runs (step * length <= 32760), so alpha of sample j is alpha0 + j * step without saturation.
// The filter lengths are fixed per sample rate (ILEN interpolation taps, TLEN + 1 tilt taps): the taps live in registers and
// the tap loops unroll.  The taps are ROM values above -32768, so L_mult's single saturating case (-32768 x -32768) cannot
// occur and the product is a plain multiply-and-double; the accumulations keep the reference's saturating adds.
template <int ILEN, int TLEN>
LC3_HD void ltpf_synth_n(i16 *out, const i16 *in, int length, int pitch_int, int pitch_fr, i32 gain, int sf, int fs_idx, int fade)
{
    i32 ifl[ILEN], tfl[TLEN + 1];
#pragma unroll
    for (int q = 0; q < ILEN; q++) ifl[q] = rom_ltpf_dec_inter[(fs_idx * 4 + pitch_fr) * 12 + q];
#pragma unroll
    for (int q = 0; q <= TLEN; q++) tfl[q] = rom_ltpf_dec_tilt[(fs_idx * 4 + sf) * 11 + q];
    const int x0off = -pitch_int + rom_ltpf_dec_inter_shift[fs_idx];
    i32 alpha0 = 0, step = 0;
    if (fade != 0) {
        if (fade < 0) alpha0 = 0x7FFF;
        step = length == 20 ? 1638 : (length == 40 ? 819 : (length == 60 ? 546 : (length == 80 ? 409 : (length == 120 ? 273 : 0))));
        if (fade < 0) step = -step;
    }
    const int C = fs_idx == 0 ? 16 : 32;
    for (int j0 = 0; j0 < length; j0 += C) {
        LC3_LANES(l)
            const int j = j0 + l;
            if (l < C && j < length) {
                const i16 *x0 = out + j + x0off, *y0 = in + j;
                i32 s = (x0[0] * ifl[0]) * 2;
#pragma unroll
                for (int q = 1; q < ILEN; q++) s = l_add(s, (x0[-q] * ifl[q]) * 2);
#pragma unroll
                for (int q = 0; q <= TLEN; q++) s = l_sub(s, (y0[-q] * tfl[q]) * 2);
                const i32 v = l_round(s);
                i32 k = s_mult_r(gain, v);
                if (fade != 0) k = s_mult_r(k, alpha0 + j * step);
                out[j] = (i16)s_add(in[j], k);
            }
        LC3_LANES_END
    }
}
LC3_HD void ltpf_synth(i16 *out, const i16 *in, int length, int pitch_int, int pitch_fr, i32 gain, int sf, int fs_idx, int fade)
{ // inter_filter_len / tilt_filter_len of fx/constants.c:2497, :2514
    switch (fs_idx) {
    case 0: case 1: ltpf_synth_n<4, 2>(out, in, length, pitch_int, pitch_fr, gain, sf, fs_idx, fade); break;
    case 2: ltpf_synth_n<6, 4>(out, in, length, pitch_int, pitch_fr, gain, sf, fs_idx, fade); break;
    case 3: ltpf_synth_n<8, 6>(out, in, length, pitch_int, pitch_fr, gain, sf, fs_idx, fade); break;
    default: ltpf_synth_n<12, 10>(out, in, length, pitch_int, pitch_fr, gain, sf, fs_idx, fade); break;
    }
}

// yrow: DCT-IV outputs, row (f * n_rows + li) of N words.  win: the low-delay window of this configuration (&rom_win[b.W_off];
// the kernel passes a copy in shared memory, its indices differ from lane to lane).
template <int BITS> LC3_HD void synth_warp_body(const LC3D_Batch &b, int li, SynthShared &sh, const i32 *yrow, const i16 *win)
{
    if (li >= b.n_lanes) return;
    const int tile = li >> 5, tl = li & 31, slot = b.slot[li];
    SV<i32> st = slot_view(b.st32, slot, DST32_N);
    SV<i16> s16 = slot_view(b.st16, slot, DST16_N);
    SV<const i32> cfg = tile_view(b.cfg, (size_t)tile, DCFG_N, tl);
    const int N = b.N, fs_idx = b.fs_idx, ola_len = b.ola_len, x_len = b.ltpf_x_len, y_len = b.ltpf_y_len;
    const int z = b.la_zeroes, m = N >> 1, o = m - z;
    const int N4 = rom_ltpf_dec_overlap[fs_idx], N34 = N - N4, tilt_len = rom_ltpf_dec_tilt_len[fs_idx];
    const i32 sf_idx = cfg[DCFG_LTPF_SCALE_IDX];

    LC3_LANES(l)
        for (int k = l; k < ola_len; k += 32) sh.ola[k] = s16[DS16_OLA + k];
        for (int k = l; k < x_len; k += 32) sh.bx[k] = s16[DS16_LTPF_X + k];
        for (int k = l; k < y_len; k += 32) sh.by[k] = s16[DS16_LTPF_Y + k];
    LC3_LANES_END
    i32 mem_e = st[DST_OLA_EXP], old_e = st[DST_LTPF_E], old_pitch_int = st[DST_LTPF_PITCH_INT], old_pitch_fr = st[DST_LTPF_PITCH_FR];
    i32 old_gain = st[DST_LTPF_GAIN], mem_active = st[DST_LTPF_ACTIVE], old_sf_idx = st[DST_LTPF_SCALE_IDX];

    for (int f = 0; f < b.n_frames; f++) {
        SV<const i32> dout = tile_view((const i32 *)b.dout, dframe_tile(b, f, tile), DO_N, tl);
        const i32 *y = yrow + dframe_row(b, f, li) * N;
        i32 y_e = dout[DO_Y_E], y_s = dout[DO_Y_S];

        // ---- windowing + overlap-add (fx/imdct_fx.c:105-166)
        i32 mem_s = warp_headroom16(sh.ola, ola_len, 1, sh.red);
        if (mem_s < 16) { mem_s = s_sub(mem_s, 1); mem_e = s_sub(mem_e, mem_s); }
        else mem_e = y_e;
        const i32 sd = s_sub(mem_e, y_e);
        if (sd > 0) { y_s = s_sub(y_s, sd); y_e = s_add(y_e, sd); }
        else { mem_s = s_add(mem_s, sd); mem_e = s_sub(mem_e, sd); }
        mem_s = imax(mem_s, -31); y_s = imax(y_s, -31);
        // the two shift counts are the same for the whole frame: split once (l_shl_plan == l_shl for every value and count,
        // tests/hostsim/primcheck.cpp); the window ROM holds no -32768 (tests/test_rom_bounds.py), so Mpy_32_16 cannot saturate
        const ShlPlan pm = shl_plan(mem_s), py = shl_plan(y_s);
        LC3_LANES(l)
            for (int i = l; i < o; i += 32)
                sh.xf[i] = (i16)l_round(l_sub(l_shl_plan(l_dep_h(sh.ola[i]), pm), mpy32x16k(l_shl_plan(y[m + i + z], py), win[4 * m - 1 - i - z])));
            for (int i = l; i < m; i += 32) {
                sh.xf[i + o] = (i16)l_round(l_add(l_shl_plan(l_dep_h(sh.ola[i + o]), pm), mpy32x16k(l_shl_plan(y[2 * m - 1 - i], py), win[3 * m - 1 - i])));
                const i32 v = l_shl_plan(y[i], py);
                sh.xf[3 * m - z + i] = (i16)l_round(l_neg(mpy32x16k(v, win[m - 1 - i])));
                sh.xf[3 * m - z - 1 - i] = (i16)l_round(l_neg(mpy32x16k(v, win[m + i])));
            }
        LC3_LANES_END
        LC3_LANES(l)
            for (int k = l; k < ola_len; k += 32) sh.ola[k] = sh.xf[N + k];
        LC3_LANES_END
        mem_e = y_e;
        i32 x_e = y_e;

        // ---- LTPF (fx/ltpf_decoder_fx.c:57-262, bfi = 0)
        const i32 ltpf = dout[DO_LTPF0], pitch_index = dout[DO_LTPF2];
        i32 ltpf_active = dout[DO_LTPF1], pitch_int = 0, pitch_fr = 0, gain = 0;
        if (ltpf != 0) {
            if (pitch_index < 380) { pitch_int = (pitch_index + 64) >> 2; pitch_fr = pitch_index - (pitch_int << 2) + 128; }
            else if (pitch_index < 440) { pitch_int = (pitch_index - 126) >> 1; pitch_fr = (pitch_index << 1) - (pitch_int << 2) - 252; }
            else { pitch_int = pitch_index - 283; pitch_fr = 0; }
            i32 pitch = (pitch_int << 2) + pitch_fr;
            pitch = s_mult_r(pitch << 2, (i32)rom_ltpf_dec_pitch_scale[fs_idx]);
            pitch_int = pitch >> 2;
            pitch_fr = pitch - (pitch_int << 2);
        }
        if (sf_idx < 0) ltpf_active = 0;
        if (ltpf_active != 0) gain = rom_ltpf_dec_gain[sf_idx];

        i16 *by = sh.by;
        if (ltpf_active == 0 && mem_active == 0) {
            const i32 s = s_sub(old_e, x_e);
            const int keep = y_len - N;
            if (s > 0) {
                warp_shift_down(by, keep, N);
                LC3_LANES(l)
                    if (s > 15) {
                        for (int k = l; k < N; k += 32) by[keep + k] = 0;
                        for (int k = l; k < x_len; k += 32) sh.bx[k] = 0;
                    } else {
                        for (int k = l; k < N; k += 32) by[keep + k] = (i16)s_shr(sh.xf[k], s);
                        for (int k = l; k < x_len; k += 32) sh.bx[k] = (i16)s_shr(sh.xf[k + N - x_len], s);
                    }
                LC3_LANES_END
            } else {
                warp_shift_down(by, keep, N);
                LC3_LANES(l)
                    if (s < -15) { for (int k = l; k < keep; k += 32) by[k] = 0; }
                    else if (s < 0) { for (int k = l; k < keep; k += 32) by[k] = (i16)s_shl(by[k], s); }
                    for (int k = l; k < N; k += 32) by[keep + k] = sh.xf[k];
                    for (int k = l; k < x_len; k += 32) sh.bx[k] = sh.xf[k + N - x_len];
                LC3_LANES_END
                old_e = x_e;
            }
            old_pitch_int = pitch_int; old_pitch_fr = pitch_fr; old_gain = 0; mem_active = 0;
        } else {
            i16 *x = sh.bx + x_len, *yo = by + y_len;
            LC3_LANES(l)
                for (int k = l; k < N; k += 32) x[k] = sh.xf[k];
            LC3_LANES_END
            const i32 s0 = s_sub(imin(warp_headroom16(sh.bx, x_len, 1, sh.red), warp_headroom16(by, y_len, 1, sh.red)), 1);
            old_e = s_sub(old_e, s0);
            const i32 s1 = s_sub(warp_headroom16(x, N, 0, sh.red), 1);
            x_e = s_sub(x_e, s1);
            const i32 s = s_sub(old_e, x_e);
            if (s > 0) {
                warp_scale_sig(x, N, s_sub(s1, s));
                warp_scale_sig(sh.bx, x_len, s0);
                warp_scale_sig(by, y_len, s0);
                x_e = old_e;
            } else {
                warp_scale_sig(x, N, s1);
                warp_scale_sig(sh.bx, x_len, s_add(s0, s));
                warp_scale_sig(by, y_len, s_add(s0, s));
                old_e = x_e;
            }
            if (ltpf_active == 0) {
                ltpf_synth(yo, x, N4, old_pitch_int, old_pitch_fr, old_gain, old_sf_idx, fs_idx, -1);
            } else if (mem_active == 0) {
                ltpf_synth(yo, x, N4, pitch_int, pitch_fr, gain, sf_idx, fs_idx, 1);
            } else if (pitch_int == old_pitch_int && old_pitch_fr == pitch_fr) {
                ltpf_synth(yo, x, N4, pitch_int, pitch_fr, gain, sf_idx, fs_idx, 0);
            } else {
                ltpf_synth(yo, x, N4, old_pitch_int, old_pitch_fr, old_gain, old_sf_idx, fs_idx, -1);
                LC3_LANES(l)
                    for (int k = l; k < N4 + tilt_len; k += 32) sh.z[k] = yo[k - tilt_len];
                LC3_LANES_END
                ltpf_synth(yo, sh.z + tilt_len, N4, pitch_int, pitch_fr, gain, sf_idx, fs_idx, 1);
            }
            if (ltpf_active > 0) {
                ltpf_synth(yo + N4, x + N4, N34, pitch_int, pitch_fr, gain, sf_idx, fs_idx, 0);
            } else {
                LC3_LANES(l)
                    for (int k = N4 + l; k < N; k += 32) yo[k] = x[k];
                LC3_LANES_END
            }
            // the frame's output goes back to xf (the PCM is written from there); then the memories move on
            LC3_LANES(l)
                for (int k = l; k < N; k += 32) sh.xf[k] = yo[k];
            LC3_LANES_END
            warp_shift_down(by, y_len, N);
            warp_shift_down(sh.bx, x_len, N);
            old_pitch_int = pitch_int; old_pitch_fr = pitch_fr; old_gain = gain; mem_active = ltpf_active;
        }
        old_sf_idx = sf_idx;

        // ---- output scaling (fx/dec_lc3.c:289-311)
        if (BITS == 16) {
            const ShlPlan po = shl_plan(-s_sub(15, x_e)); // L_shr by scale = L_shl by -scale, the same for the whole frame
            i16 *dst = (i16 *)b.pcm + b.pcm_off[li] + (size_t)(b.frame0 + f) * N;
            LC3_LANES(l)
                for (int k = l; k < N; k += 32) dst[k] = (i16)l_round(l_shl_plan(l_dep_h(sh.xf[k]), po));
            LC3_LANES_END
        } else {
            const i32 scale = s_sub(s_sub(31 + 16, 24), x_e);
            const i32 offset = l_shr(32768, s_sub(16, scale));
            const ShlPlan po = shl_plan(-scale);
            i32 *dst = (i32 *)b.pcm + b.pcm_off[li] + (size_t)(b.frame0 + f) * N;
            LC3_LANES(l)
                for (int k = l; k < N; k += 32) dst[k] = l_shl_plan(l_add(l_dep_h(sh.xf[k]), offset), po);
            LC3_LANES_END
        }
    }

    LC3_LANES(l)
        for (int k = l; k < ola_len; k += 32) s16[DS16_OLA + k] = sh.ola[k];
        for (int k = l; k < x_len; k += 32) s16[DS16_LTPF_X + k] = sh.bx[k];
        for (int k = l; k < y_len; k += 32) s16[DS16_LTPF_Y + k] = sh.by[k];
    LC3_LANES_END
    LC3_LANE0
        st[DST_OLA_EXP] = mem_e; st[DST_LTPF_E] = old_e; st[DST_LTPF_PITCH_INT] = old_pitch_int; st[DST_LTPF_PITCH_FR] = old_pitch_fr;
        st[DST_LTPF_GAIN] = old_gain; st[DST_LTPF_ACTIVE] = mem_active; st[DST_LTPF_SCALE_IDX] = old_sf_idx;
    LC3_LANE0_END
}

} // namespace lc3k
