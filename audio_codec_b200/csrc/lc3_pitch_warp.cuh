// Time-domain pass, ONE WARP PER CHANNEL-STREAM (fx/enc_lc3.c:65-114): [24-bit block scaling
// fx/scale_signal24_fx.c:13-92] 12.8 kHz resampler + 50 Hz high-pass (fx/resamp12k8_fx.c:15-113), open-loop
// pitch analysis (fx/olpa_fx.c:16-222), LTPF parameter coder (fx/ltpf_coder_fx.c:17-307), attack detector
// (fx/attack_detector_fx.c:13-119).
//
// This is the chain whose state makes frame n+1 wait for frame n, so a stream's frames are walked in order;
// the parallelism is inside the frame: lanes split the FIR outputs, the correlation lags and the sample
// loops, with the stream's signal memories resident in the warp's shared-memory block for the whole chunk.
// Phases and reductions follow lc3_warp.cuh; arithmetic follows the cited reference lines exactly.  Sums of
// products that the reference accumulates with saturating L_mac0 are reduced across lanes in 64 bits and
// clamped: operands are block-scaled to |v| <= 2^12 (3 bits of headroom, olpa_fx.c:60-98,
// ltpf_coder_fx.c:63-79), so no partial sum of <= 256 products can leave the 32-bit range and the result is
// the same for any summation order.
#pragma once
#include "lc3_stages.cuh"
#include "lc3_warp.cuh"

namespace lc3k {

struct PitchShared {
    WarpRed red;
    i32 sum0[20], sum2[20];       // LTPF: cross sums and lagged energies per candidate lag
    i16 xs[64 + LC3K_MAX_N];      // scaled input; xs[64 + i] = sample i of the frame, i in [-60, N)
    i16 filt[240];                // polyphase resampler taps of this sample rate
    i16 s12[24 + 128 + 8];        // [24-sample delay line | new 12.8 kHz samples]
    i16 fir[128];                 // resampler output before the high-pass
    i16 m12[4];                   // olpa_mem_s12k8
    i16 olpa[2][4 + 130 + 64 + 2]; // 6.4 kHz history + new samples, ping-pong; 4 samples of pad in front (phase E reads 2 of them)
    i16 ltpf[2][4 + 264 + 128 + 4]; // 12.8 kHz history + new samples, ping-pong; 4 samples of pad in front (lag windows read 3)
    i16 ac[32];                   // LTPF normalised correlations
    i32 xc[224];                  // OLPA autocorrelation partial sums [7 lags][lane], then LTPF cross-correlation partial sums [lane][4 lags]
    i16 dd[2 + 160 + 2];          // attack detector: decimated signal, two samples of history first
};

// raw sample i (relative to frame f) of this stream: previous-frame samples of the chunk's first frame come
// from the staged history prefix, everything else straight from the caller's PCM (coalesced along i)
template <class T> struct RawPcm {
    const T *cur;        // caller's PCM at sample 0 of frame f
    SV<const T> hist;    // staged history frame (tile-interleaved), sample 0 of the previous frame
    int N, first;
    LC3_HDM i32 operator()(int i) const { return (i < 0 && first) ? (i32)hist[N + i] : (i32)cur[i]; }
};

// OLPA autocorrelation, lags 17..114 (olpa_fx.c:101-122): lane = (group of 7 consecutive lags, half of the samples).
// 14 groups x 2 halves = 28 lanes hold 7 accumulators each; a lane walks SEG samples with a seven-sample window of the lagged
// signal sliding through registers (everything unrolled: one load per new window sample, no register rotation).  A half has
// at most 32 products of |v| <= 2^12 operands per lag, the two halves are added afterwards: far below 2^31, any order.
template <int SEG> LC3_HD void olpa_acf_half(const i16 *s, int l, i32 *xc /* [7][32] */)
{
    const int g = l >> 1, half = l & 1;
    i32 a[7] = {0, 0, 0, 0, 0, 0, 0};
    if (g < 14) {
        const i16 *x = s + half * SEG;   // x[m] = s[half * SEG + m]
        const i16 *q = x - (17 + 7 * g); // q[m - j] = the sample lag (17 + 7 g + j) behind x[m]
        i32 w[SEG + 6];                  // w[k] = q[k - 6]
#pragma unroll
        for (int k = 0; k < SEG + 6; k++) w[k] = q[k - 6];
#pragma unroll
        for (int m = 0; m < SEG; m++) {
            const i32 xv = x[m];
#pragma unroll
            for (int j = 0; j < 7; j++) a[j] += xv * w[m - j + 6];
        }
    }
#pragma unroll
    for (int j = 0; j < 7; j++) xc[j * 32 + l] = a[j];
}

// PRE = true (16-bit input only): the resampler FIR, the high-pass recursion and the attack detector have already
// run in their own passes (fir_warp_body, scalar_body below) and b.r12 holds the high-passed 12.8 kHz samples;
// this body then does the pitch analysis only.  PRE = false: everything in one pass (24-bit input, whose block
// exponent depends on the filter state and therefore cannot be split).
// sync_threads > 0 (kernel only): the block's active warps meet at a barrier before every frame, so that they walk the
// (instruction-cache sized) body together and share its fetches.
#if defined(__CUDA_ARCH__)
#define LC3_PITCH_SYNC() do { if (sync_threads > 0) asm volatile("bar.sync 1, %0;" ::"r"(sync_threads) : "memory"); } while (0)
#else
#define LC3_PITCH_SYNC() do { } while (0)
#endif
// DMS = 25 / 50 / 100 fixes the frame duration at compile time (0: taken from the batch): every lane-strided loop of the body
// then has a constant trip count and unrolls without the remainder bookkeeping a run-time bound needs.
template <int BITS, bool PRE, int DMS = 0> LC3_HD void pitch_warp_body(const LC3K_Batch &b, int li, PitchShared &sh, int sync_threads = 0)
{
    typedef typename std::conditional<BITS == 16, i16, i32>::type T;
    if (li >= b.n_lanes) return;
    const int tile = li >> 5, tl = li & 31;
    const int slot = b.slot[li];
    SV<i32> st = slot_view(b.st32, slot, ST32_N);
    SV<i16> s16 = slot_view(b.st16, slot, ST16_N);
    SV<const i32> cfg = slot_view(b.cfg, slot, CFG_N);
    const i32 ltpf_enable = cfg[CFG_LTPF_ENABLE], attack_handling = cfg[CFG_ATTACK];
    const int N = b.N, dms = DMS ? DMS : b.dms, rl = b.r12k8_mem_in_len, ml = b.mdct_mem_len;
    const int len = 32 * (dms / 25), len2 = len >> 1;
    const int olpa_len = dms == 25 ? 130 : 114, acflen6 = dms == 25 ? len2 + 16 : len2;
    const int lt_len = DMS ? (DMS == 25 ? 232 + 32 : 232) : b.ltpf_mem_in_len;
    const int upfac = rom_resamp_params[b.fs_idx * 4 + 0], delay2 = 2 * rom_resamp_params[b.fs_idx * 4 + 1];
    const int off_int = rom_resamp_params[b.fs_idx * 4 + 2], off_frac = rom_resamp_params[b.fs_idx * 4 + 3];

    // ---- load the stream's state: signal memories to shared memory, scalars to (replicated) registers
    LC3_LANES(l)
        for (int k = l; k < 240; k += 32) sh.filt[k] = rom_resamp_filt[b.fs_idx * 240 + k];
        if (l < 24) sh.s12[l] = s16[S16_MEM_OUT + l];
        if (l < 3) sh.m12[l] = s16[S16_OLPA_S12 + l];
        for (int k = l; k < olpa_len; k += 32) sh.olpa[0][4 + k] = s16[S16_OLPA_BUF + k];
        for (int k = l; k < lt_len; k += 32) sh.ltpf[0][4 + k] = s16[S16_LTPF_BUF + k];
    LC3_LANES_END
    i32 m0 = st[ST_MEM50_0], m1 = st[ST_MEM50_1];
    i32 olpa_exp = st[ST_OLPA_EXP], olpa_pitch = st[ST_OLPA_PITCH];
    i32 ltpf_exp = st[ST_LTPF_EXP], ltpf_nc = st[ST_LTPF_NC], ltpf_mnc = st[ST_LTPF_MNC], ltpf_on = st[ST_LTPF_ON],
        ltpf_pitch_mem = st[ST_LTPF_PITCH];
    AttackState as;
    as.filter_mem[0] = st[ST_ATT_F0]; as.filter_mem[1] = st[ST_ATT_F1]; as.detected = st[ST_ATT_DET];
    as.position = st[ST_ATT_POS]; as.scaling = st[ST_ATT_SCAL]; as.acc_energy = st[ST_ATT_ACC];
    // cleared whenever set_bitrate switched the detector off (setup_enc_lc3.c:252-275): while it is off, and -- for an off/on
    // toggle with no frame in between -- when the host moved the lane's generation word on
    const i32 att_gen = cfg[CFG_ATT_GEN];
    if (!attack_handling || att_gen != st[ST_ATT_GEN]) {
        as.filter_mem[0] = as.filter_mem[1] = 0; as.detected = as.position = as.scaling = as.acc_energy = 0;
    }
    i32 x_exp = st[ST_X_EXP], resamp_exp = st[ST_RESAMP_EXP];
    int ob = 0, lb = 0; // ping-pong selectors of the OLPA / LTPF buffers

    RawPcm<T> raw;
    raw.N = N;
    raw.hist = tile_view((const T *)b.pcm_t, (size_t)tile, (size_t)(b.n_frames + 1) * N, tl);

    for (int f = 0; f < b.n_frames; f++) {
        LC3_PITCH_SYNC();
        raw.cur = (const T *)b.pcm + b.pcm_off[li] + (size_t)(b.frame0 + f) * N;
        raw.first = f == 0;

        // ---- A. input: 24-bit block exponent (scale_signal24_fx.c:13-92), scaled samples to shared memory
        if (BITS == 24) {
            i32 e = 7;
            LC3_LANES(l) // current frame
                Range r; range_init(r);
                for (int i = l; i < N; i += 32) range_add(r, raw(i));
                sh.red.a[l] = r.mx; sh.red.b[l] = r.mn;
            LC3_LANES_END
            { Range r = {warp_max(sh.red.a), warp_min(sh.red.b)}; e = s_max(e, s_sub(23, range_headroom32(r, 32))); }
            LC3_LANES(l) // MDCT memory = last ml samples of the previous frame
                Range r; range_init(r);
                for (int i = l - ml; i < 0; i += 32) range_add(r, raw(i));
                sh.red.a[l] = r.mx; sh.red.b[l] = r.mn;
            LC3_LANES_END
            { Range r = {warp_max(sh.red.a), warp_min(sh.red.b)}; e = s_max(e, s_sub(23, range_headroom32(r, 32))); }
            LC3_LANES(l) // resampler memory = last rl samples of the previous frame
                Range r; range_init(r);
                for (int i = l - rl; i < 0; i += 32) range_add(r, raw(i));
                sh.red.a[l] = r.mx; sh.red.b[l] = r.mn;
            LC3_LANES_END
            { Range r = {warp_max(sh.red.a), warp_min(sh.red.b)}; e = s_max(e, s_sub(23, range_headroom32(r, 32))); }
            { Range r; range_init(r); range_add(r, m0); range_add(r, m1);
              e = s_max(e, s_sub(s_sub(resamp_exp, 2), range_headroom32(r, 32))); }
            LC3_LANES(l) // resampler output delay line
                Range r; range_init(r);
                if (l < 24) range_add(r, sh.s12[l]);
                sh.red.a[l] = r.mx; sh.red.b[l] = r.mn;
            LC3_LANES_END
            { Range r = {warp_max(sh.red.a), warp_min(sh.red.b)}; e = s_max(e, s_sub(s_sub(resamp_exp, 2), range_headroom16(r, 16))); }
            { Range r; range_init(r); range_add(r, sh.m12[0]); range_add(r, sh.m12[1]); range_add(r, sh.m12[2]);
              e = s_max(e, s_sub(s_sub(resamp_exp, 2), range_headroom16(r, 16))); }
            LC3_SYNC; // every lane has read m12 / s12 before they are rescaled
            x_exp = e;
            const i32 shm = s_sub(s_sub(resamp_exp, 2), x_exp);
            resamp_exp = s_add(x_exp, 2);
            if (shm) {
                m0 = l_shl(m0, shm); m1 = l_shl(m1, shm);
                LC3_LANES(l)
                    if (l < 24) sh.s12[l] = (i16)s_shl(sh.s12[l], shm);
                    if (l >= 24 && l < 27) sh.m12[l - 24] = (i16)s_shl(sh.m12[l - 24], shm);
                LC3_LANES_END
            }
        }
        if (PRE) {
            SV<const i16> r = tile_view((const i16 *)b.r12, frame_tile(b, f, tile), 128, tl);
            LC3_LANES(l)
                for (int n = l; n < len; n += 32) sh.s12[24 + n] = r[n];
                if (attack_handling) for (int i = l; i < N; i += 32) sh.xs[64 + i] = (i16)raw(i); // for the attack detector
            LC3_LANES_END
        } else {
            {
                const i32 shx = s_sub(23, x_exp);
                LC3_LANES(l)
                    for (int i = l - rl; i < N; i += 32) sh.xs[64 + i] = (i16)(BITS == 16 ? raw(i) : l_round(l_shl(raw(i), shx)));
                LC3_LANES_END
            }

            // ---- B. polyphase FIR to 12.8 kHz (resamp12k8_fx.c:62-86): lane = output sample
            LC3_LANES(l)
                for (int n = l; n < len; n += 32) {
                    const int t = n * off_frac, ii = 1 + n * off_int + t / upfac - rl, fr = t % upfac;
                    const i16 *co = &sh.filt[fr * delay2], *in = &sh.xs[64 + ii];
                    // the reference's saturating L_mac0 chain cannot saturate: sum |taps| <= 56970 < 2^16 for every
                    // phase of every rate (checked against the ROM in tests/test_rom_bounds.py), |x| <= 2^15
                    i32 acc = 0;
                    for (int m = 0; m < delay2; m++) acc += in[m] * co[m];
                    sh.fir[n] = (i16)l_round(acc);
                }
            LC3_LANES_END

            // ---- C. 50 Hz high-pass, a true recursion (resamp12k8_fx.c:88-104): every lane runs it redundantly so
            //         the filter state stays warp-uniform; all lanes store the same value to s12[24 + n]
            {
                const i32 n0 = rom_hp50_num[0], n1 = rom_hp50_num[1], n2 = rom_hp50_num[2], d0 = rom_hp50_den[0], d1 = rom_hp50_den[1];
                for (int n = 0; n < len; n++) {
                    const i32 v = sh.fir[n];
                    const i32 o = l_mac0(m0, n0, v);
                    const i32 t = l_mac0(mpy32x16k(o, d0), n1, v);
                    m0 = l_add(m1, l_shl_p(t, 1));
                    m1 = l_mac0(mpy32x16k(o, d1), n2, v);
                    sh.s12[24 + n] = (i16)l_round(o);
                }
            }
            LC3_SYNC; // s12[0 .. len] is the delayed 12.8 kHz signal of this frame
        }

        // ---- D. OLPA: 2:1 decimation with a 5-tap low-pass (olpa_fx.c:36-58): lane = 6.4 kHz sample.
        //         v[j] = m12[3 + j] for j < 0, s12[j] otherwise; output k uses v[2k-3 .. 2k+1].
        i16 *omem = sh.olpa[ob] + 4, *s6 = omem + olpa_len;
        LC3_LANES(l)
            for (int k = l; k < len2; k += 32) {
                i32 v[5];
#pragma unroll
                for (int j = 0; j < 5; j++) { const int q = 2 * k - 3 + j; v[j] = q < 0 ? sh.m12[3 + q] : sh.s12[q]; }
                i32 sum = l_mac(l_mac(l_mult(v[0], 4053), v[1], 7712), v[2], 9239);
                s6[k] = (i16)l_round(l_mac(l_mac(sum, v[3], 7712), v[4], 4053));
            }
        LC3_LANES_END
        LC3_LANES(l)
            if (l < 3) sh.m12[l] = sh.s12[len - 3 + l];
            // headroom of the history and of the new samples (olpa_fx.c:60-63)
            Range rh, rn; range_init(rh); range_init(rn);
            for (int k = l; k < olpa_len; k += 32) range_add(rh, omem[k]);
            for (int k = l; k < len2; k += 32) range_add(rn, s6[k]);
            sh.red.a[l] = rh.mx; sh.red.b[l] = rh.mn; sh.red.c[l] = rn.mx; sh.red.d[l] = rn.mn;
        LC3_LANES_END
        i32 pitch, normcorr;
        {
            Range rh = {warp_max(sh.red.a), warp_min(sh.red.b)};
            Range rn = {warp_max(sh.red.c), warp_min(sh.red.d)};
            const i32 sc0 = s_sub(range_headroom16(rh, 16), 3);
            olpa_exp = s_sub(olpa_exp, sc0);
            const i32 sc1 = s_sub(range_headroom16(rn, 16), 3);
            const i32 s6_exp = s_sub(resamp_exp, sc1);
            const i32 sc2 = s_sub(olpa_exp, s6_exp);
            i32 e_new, e_hist;
            if (sc2 > 0) { e_new = s_sub(sc1, sc2); e_hist = sc0; }
            else { e_new = sc1; e_hist = s_add(sc0, sc2); olpa_exp = s6_exp; }
            const i32 t_new = e_new < 0 ? s_shl(-32768, s_max(e_new, -15)) : 0;
            const i32 t_hist = e_hist < 0 ? s_shl(-32768, s_max(e_hist, -15)) : 0;
            LC3_LANES(l) // Scale_sig (basop_util.c:571-595) on both parts
                if (e_new) for (int k = l; k < len2; k += 32) s6[k] = (i16)scale_val(s6[k], e_new, t_new);
                if (e_hist) for (int k = l; k < olpa_len; k += 32) omem[k] = (i16)scale_val(omem[k], e_hist, t_hist);
            LC3_LANES_END
        }
        {
            // ---- E. autocorrelation, lags 17..114: lane = lag (olpa_fx.c:101-122, :157-172)
            const i16 *s = dms == 25 ? s6 - 16 : s6;
            const i32 lo = s_max(17, s_sub(olpa_pitch, 4)), hi = s_min(114, s_add(olpa_pitch, 4));
            LC3_LANES(l)
                if (acflen6 == 64) olpa_acf_half<32>(s, l, sh.xc);
                else if (acflen6 == 48) olpa_acf_half<24>(s, l, sh.xc);
                else olpa_acf_half<16>(s, l, sh.xc);
            LC3_LANES_END
            LC3_LANES(l) // totals of the pair's two halves; lane (g, half) weighs lags j = half, half + 2, .. of its group
                i64 kw = LC3_I32_MIN * (i64)4294967296LL, k2 = kw;
                const int g = l >> 1, half = l & 1;
                if (g < 14) {
#pragma unroll
                    for (int j = 0; j < 7; j++) {
                        if ((j & 1) != half) continue;
                        const int n = 17 + 7 * g + j;
                        const i32 acc = sh.xc[j * 32 + l] + sh.xc[j * 32 + (l ^ 1)];
                        const i64 a = argmax_key(mpy32x16(acc, rom_olpa_weight[n - 17]), n);
                        kw = a > kw ? a : kw;
                        if (n >= lo && n <= hi) { const i64 c = argmax_key(acc, n); k2 = c > k2 ? c : k2; }
                    }
                }
                sh.red.q[l] = kw;
                sh.red.q2[l] = k2;
            LC3_LANES_END
            i32 p = argmax_index(warp_max64(sh.red.q));
            const i32 p2 = lo > hi ? lo : argmax_index(warp_max64(sh.red.q2));

            // ---- F. normalised correlation at the candidates (olpa_fx.c:126-155, :174-205): lane = sample
            // both candidates in one sweep over the samples; the two normalisations (a serial chain of ~190 operations each)
            // then run side by side on lanes 0 and 1 instead of one after the other on the whole warp
            i32 nc[2] = {0, 0};
            const int ncand = p != p2 ? 2 : 1;
            LC3_LANES(l)
                i32 a0 = 0, a1 = 0, a2 = 0, b0 = 0, b1 = 0;
                for (int m = l; m < acflen6; m += 32) {
                    const i32 x = s[m], y = s[m - p], z = s[m - p2];
                    a0 += x * y; a1 += y * y; a2 += x * x; b0 += x * z; b1 += z * z;
                }
                sh.red.a[l] = a0; sh.red.b[l] = a1; sh.red.c[l] = a2; sh.red.d[l] = b0; sh.xc[l] = b1;
            LC3_LANES_END
            {
                const i32 s0 = warp_sum(sh.red.a), s1 = warp_sum(sh.red.b) + 1, s2 = warp_sum(sh.red.c) + 1;
                const i32 t0 = warp_sum(sh.red.d), t1 = warp_sum(sh.xc) + 1;
                LC3_LANES(l)
                    if (l < ncand) sh.red.a[l] = normcorr_q15(l == 0 ? s0 : t0, l == 0 ? s1 : t1, s2, 0, 0); // one copy of the code for both lanes
                LC3_LANES_END
                nc[0] = sh.red.a[0];
                if (ncand == 2) nc[1] = sh.red.a[1];
                LC3_SYNC;
            }
            normcorr = nc[0];
            if (p != p2 && s_sub(nc[1], s_mult_r(nc[0], 27853)) > 0) { p = p2; normcorr = nc[1]; }
            olpa_pitch = p;
            pitch = s_shl(p, 1);
            // history for the next frame goes to the other buffer
            i16 *onext = sh.olpa[ob ^ 1] + 4;
            LC3_LANES(l)
                for (int k = l; k < olpa_len; k += 32) onext[k] = omem[k + len2];
            LC3_LANES_END
            ob ^= 1;
        }

        // ---- G. LTPF (ltpf_coder_fx.c:17-307)
        i32 ltpf_bits, lt0 = 0, lt1 = 0, lt2 = 0, nc_ltpf, pitch_ltpf = 0, active = 0;
        {
            i16 *lmem = sh.ltpf[lb] + 4, *x = lmem + lt_len;
            LC3_LANES(l)
                for (int k = l; k <= len; k += 32) x[k] = sh.s12[k];
            LC3_LANES_END
            LC3_LANES(l)
                Range rh, rn; range_init(rh); range_init(rn);
                for (int k = l; k < lt_len; k += 32) range_add(rh, lmem[k]);
                for (int k = l; k <= len; k += 32) range_add(rn, x[k]);
                sh.red.a[l] = rh.mx; sh.red.b[l] = rh.mn; sh.red.c[l] = rn.mx; sh.red.d[l] = rn.mn;
            LC3_LANES_END
            {
                Range rh = {warp_max(sh.red.a), warp_min(sh.red.b)};
                Range rn = {warp_max(sh.red.c), warp_min(sh.red.d)};
                const i32 sc0 = s_sub(range_headroom16(rh, 16), 3);
                ltpf_exp = s_sub(ltpf_exp, sc0);
                const i32 sc1 = s_sub(range_headroom16(rn, 16), 3);
                const i32 xe = s_sub(resamp_exp, sc1);
                const i32 sc2 = s_sub(ltpf_exp, xe);
                i32 e_new, e_hist;
                if (sc2 > 0) { e_new = s_sub(sc1, sc2); e_hist = sc0; }
                else { e_new = sc1; e_hist = s_add(sc0, sc2); ltpf_exp = xe; }
                const i32 t_new = e_new < 0 ? s_shl(-32768, s_max(e_new, -15)) : 0;
                const i32 t_hist = e_hist < 0 ? s_shl(-32768, s_max(e_hist, -15)) : 0;
                LC3_LANES(l)
                    if (e_new) for (int k = l; k <= len; k += 32) x[k] = (i16)scale_val(x[k], e_new, t_new);
                    if (e_hist) for (int k = l; k < lt_len; k += 32) lmem[k] = (i16)scale_val(lmem[k], e_hist, t_hist);
                LC3_LANES_END
            }
            if (s_sub(normcorr, 19660) > 0) {
                const i32 min_p = s_max(s_sub(pitch, 4), 32), max_p = s_min(s_add(pitch, 4), 228);
                const i32 ac_lo = s_sub(min_p, 4), ac_hi = s_add(max_p, 4);
                const int nlags = ac_hi - ac_lo + 1;
                const int acflen = dms == 25 ? 2 * len : len;
                const i16 *xp = dms == 25 ? x - len : x;

                LC3_LANES(l) // energies of the current and of the first lagged window: lane = sample
                    i64 a1 = 0, a2 = 0;
                    for (int m = l; m < acflen; m += 32) { i32 u = xp[m], v = xp[m - ac_lo]; a1 += u * u; a2 += v * v; }
                    sh.red.q[l] = a1; sh.red.q2[l] = a2;
                LC3_LANES_END
                i32 sum1 = clamp64(1 + warp_sum64(sh.red.q));
                const i32 sum2_0 = clamp64(1 + warp_sum64(sh.red.q2));
                const i32 n1 = l_norm(sum1);
                sum1 = l_shl(sum1, n1);

                // cross-correlations of the <= 17 candidate lags: lane = (four consecutive lags, one sixth of the samples),
                // a four-sample window of the lagged signal slides through registers.  A segment has <= 22 products of
                // |v| <= 2^12 operands, so its partial sums fit 32 bits; the six partials are added exactly below.
                const int seg_len = (acflen + 5) / 6;
                LC3_LANES(l)
                    const int qd = l / 6, sg = l - 6 * qd;
                    i32 a0 = 0, a1 = 0, a2 = 0, a3 = 0;
                    i64 e_all = 0;
                    if (qd < 5) {
                        const int m_lo = sg * seg_len, m_hi = imin(acflen, m_lo + seg_len);
                        const i16 *q = xp - (ac_lo + 4 * qd); // q[m - j] = xp[m - (ac_lo + 4 qd + j)]
                        i32 w1 = q[m_lo - 1], w2 = q[m_lo - 2], w3 = q[m_lo - 3];
                        for (int m = m_lo; m < m_hi; m++) {
                            const i32 x = xp[m], w0 = q[m];
                            a0 += x * w0; a1 += x * w1; a2 += x * w2; a3 += x * w3;
                            w3 = w2; w2 = w1; w1 = w0;
                        }
                    }
                    sh.xc[4 * l] = a0; sh.xc[4 * l + 1] = a1; sh.xc[4 * l + 2] = a2; sh.xc[4 * l + 3] = a3;
                    // energy of every sample any lagged window touches (bounds all window energies from above)
                    for (int m = -ac_hi + l; m < acflen - ac_lo; m += 32) { const i32 v = xp[m]; e_all += v * v; }
                    sh.red.q[l] = e_all;
                    // change of the lagged window's energy from lag n-1 to lag n = ac_lo + l (ltpf_coder_fx.c:116-120)
                    i32 dlt = 0;
                    if (l >= 1 && l < nlags) { const i32 o = xp[acflen - (ac_lo + l)], w = xp[-(ac_lo + l)]; dlt = w * w - o * o; }
                    sh.red.a[l] = dlt;
                LC3_LANES_END
                // If even the union of the windows has an energy below 2^31, no window energy and no intermediate of the
                // reference's L_msu0/L_mac0 chain can saturate, and the chain is a plain prefix sum.
                const bool plain = 1 + warp_sum64(sh.red.q) <= (i64)LC3_I32_MAX;
                LC3_LANES(l) // lane = candidate lag
                    if (l < nlags) {
                        const int n = ac_lo + l, qd = l >> 2, j = l & 3;
                        i64 s64 = 0;
                        for (int sg = 0; sg < 6; sg++) s64 += sh.xc[4 * (6 * qd + sg) + j];
                        const i32 s0 = clamp64(s64);
                        i32 s2 = sum2_0;
                        if (plain) {
                            for (int k = 1; k <= l; k++) s2 += sh.red.a[k];
                        } else {
                            for (int q = ac_lo + 1; q <= n; q++) {
                                const i32 o = xp[acflen - 1 - (q - 1)], w = xp[-q];
                                s2 = l_msu0(s2, o, o);
                                s2 = l_mac0(s2, w, w);
                            }
                        }
                        sh.ac[l] = (i16)normcorr_q15(s0, sum1, s2, 1, n1);
                    }
                LC3_LANES_END

                i32 ac_max = sh.ac[min_p - ac_lo], pit = min_p;
                for (int n = min_p + 1; n <= max_p; n++) {
                    const i32 v = sh.ac[n - ac_lo];
                    if (s_sub(v, ac_max) > 0) pit = n;
                    ac_max = s_max(ac_max, v);
                }
                i32 p_int = pit, p_fr = 0, p_idx = s_add(p_int, 283);
                if (pit < 157) { // fractional refinement
                    const int res = pit < 127 ? 1 : 2;
                    int fr0 = pit < 127 ? -3 : -2;
                    if (pit == min_p) fr0 = 0;
                    // one candidate fraction per lane (at most 7), first strict maximum wins as in the reference's scan
                    LC3_LANES(l)
                        const int fr = fr0 + l * res;
                        i64 key = LC3_I32_MIN * (i64)4294967296LL;
                        if (fr < 4) {
                            const i16 *h = &rom_ltpf_ac_interp[(fr + 3) * 9];
                            i32 s = l_mult0(sh.ac[p_int - ac_lo - 4], h[0]);
                            for (int k = 1; k < 9; k++) s = l_mac0(s, sh.ac[p_int - ac_lo - 4 + k], h[k]);
                            if (s != LC3_I32_MIN) key = argmax_key(s, l); // a sum of -2^31 never beats the scan's initial value
                        }
                        sh.red.q[l] = key;
                    LC3_LANES_END
                    const i64 bk = warp_max64(sh.red.q);
                    if (bk != LC3_I32_MIN * (i64)4294967296LL) p_fr = fr0 + argmax_index(bk) * res;
                    if (p_fr < 0) { p_int = s_sub(p_int, 1); p_fr = s_add(p_fr, 4); }
                    if (p_int >= 127) p_idx = s_add(s_add(s_shl(p_int, 1), p_fr >> 1), 126);
                    else p_idx = s_sub(s_add(s_shl(p_int, 2), p_fr), 128);
                }
                pitch_ltpf = s_add(s_shl(p_int, 2), p_fr);

                const i16 *f0 = &rom_ltpf_inter_filt[0], *fp = &rom_ltpf_inter_filt[p_fr * 12];
                const i32 f00 = f0[0], f01 = f0[1], f02 = f0[2], fp0 = fp[0], fp1 = fp[1], fp2 = fp[2], fp3 = fp[3];
                LC3_LANES(l) // filtered current / predicted frame and their correlation: lane = sample
                    i64 a0 = 0, a1 = 0, a2 = 0;
                    for (int n = l; n < acflen; n += 32) {
                        // |xp| <= 2^12 after the block scaling above and sum |taps| <= 32875 (tests/test_rom_bounds.py): the
                        // reference's L_mult / L_mac / round chain stays below 2^29 and is plain arithmetic
                        const i32 c = (2 * (xp[n + 1] * f00 + xp[n] * f01 + xp[n - 1] * f02) + 0x8000) >> 16;
                        const i32 p = (2 * (xp[n - p_int + 1] * fp0 + xp[n - p_int] * fp1 + xp[n - p_int - 1] * fp2 +
                                            xp[n - p_int - 2] * fp3) + 0x8000) >> 16;
                        a0 += c * p; a1 += p * p; a2 += c * c;
                    }
                    sh.red.q[l] = a0; sh.red.q2[l] = a1; sh.red.q3[l] = a2;
                LC3_LANES_END
                const i32 s0 = clamp64(warp_sum64(sh.red.q)), s1 = clamp64(1 + warp_sum64(sh.red.q2)),
                          s2 = clamp64(1 + warp_sum64(sh.red.q3));
                nc_ltpf = normcorr_q15(s0, s1, s2, 0, 0);
                if (nc_ltpf < 0) nc_ltpf = 0;

                if (ltpf_enable == 1) {
                    if ((ltpf_on == 0 && s_sub(ltpf_nc, 30802) > 0 && s_sub(nc_ltpf, 30802) > 0 &&
                         (dms == 100 || s_sub(ltpf_mnc, 30802) > 0)) ||
                        (ltpf_on == 1 && s_sub(nc_ltpf, 29491) > 0) ||
                        (ltpf_on == 1 && s_sub(s_abs(s_sub(pitch_ltpf, ltpf_pitch_mem)), 8) < 0 &&
                         s_add(s_sub(nc_ltpf, ltpf_nc), 3277) > 0 && s_sub(nc_ltpf, 27525) > 0))
                        active = 1;
                }
                lt0 = 1; lt1 = active; lt2 = p_idx;
                ltpf_bits = 11;
            } else {
                nc_ltpf = normcorr;
                ltpf_bits = 1;
            }
            i16 *lnext = sh.ltpf[lb ^ 1] + 4;
            LC3_LANES(l)
                for (int k = l; k < lt_len; k += 32) lnext[k] = lmem[k + len];
            LC3_LANES_END
            lb ^= 1;
            ltpf_mnc = ltpf_nc; ltpf_nc = nc_ltpf; ltpf_on = active; ltpf_pitch_mem = pitch_ltpf;
        }

        // ---- H. attack detector (attack_detector_fx.c:13-119), 10 ms frames at 32/48 kHz only
        if (attack_handling) {
            const i32 in_scaling = s_sub(x_exp, 15);
            LC3_LANES(l)
                Range r; range_init(r);
                for (int i = l; i < N; i += 32) range_add(r, sh.xs[64 + i]);
                sh.red.a[l] = r.mx; sh.red.b[l] = r.mn;
            LC3_LANES_END
            Range rg = {warp_max(sh.red.a), warp_min(sh.red.b)};
            const i32 s0 = s_add(range_headroom16(rg, 0), in_scaling);
            Range r1; range_init(r1); range_add(r1, as.filter_mem[0]); range_add(r1, as.filter_mem[1]);
            const i32 s1 = s_add(range_headroom16(r1, 16), as.scaling);
            Range r2; range_init(r2); range_add(r2, as.acc_energy);
            const i32 s2 = s_shr(s_add(s_add(range_headroom32(r2, 32), s_shl(as.scaling, 1)), 1), 1);
            const i32 ns = s_sub(s_min(s0, s_min(s1, s2)), 2);
            const i32 rs = s_sub(ns, as.scaling);
            if (rs) {
                as.filter_mem[0] = s_shl(as.filter_mem[0], rs);
                as.filter_mem[1] = s_shl(as.filter_mem[1], rs);
                as.acc_energy = l_shl(as.acc_energy, s_shl(rs, 1));
            }
            as.scaling = ns;
            const int dec = b.fs == 32000 ? 2 : 3;
            const i32 ds = s_sub(dec - 1, s_sub(ns, in_scaling));
            const i32 fm0 = as.filter_mem[0], fm1 = as.filter_mem[1];
            // ns <= headroom(input) + in_scaling - 2 and ns <= headroom(filter_mem) + scaling - 2, so a left shift (ds < 0) is
            // by at most headroom - dec - 1 bits and every shifted sample is within +-2^(14 - dec): the two or three
            // shr/add of the reference cannot saturate and |dd| <= 2^13 throughout (the rescaled memory included)
            const i32 lsh = 1 << imax(-ds, 0), rsh = imin(imax(ds, 0), 15);
            LC3_LANES(l) // decimate to 16 kHz: lane = output sample
                if (l == 0) { sh.dd[0] = (i16)fm0; sh.dd[1] = (i16)fm1; }
                for (int i = l; i < 160; i += 32) {
                    const i16 *in = &sh.xs[64 + dec * i];
                    i32 d = ((in[0] * lsh) >> rsh) + ((in[1] * lsh) >> rsh);
                    if (dec == 3) d += (in[2] * lsh) >> rsh;
                    sh.dd[2 + i] = (i16)d;
                }
            LC3_LANES_END
            // 3-tap high-pass on the unfiltered neighbours, energies of the four 40-sample blocks.  |dd| <= 2^13: the
            // L_mult / L_msu / L_mac chain stays within 2^29, a squared output within 2^27; a lane adds at most two of them
            // to a block (samples l, l + 32, ..: block i / 40 of pass `it` is one of two known at compile time)
            LC3_LANES(l)
                i32 e[4] = {0, 0, 0, 0};
#pragma unroll
                for (int it = 0; it < 5; it++) {
                    const int i = l + 32 * it;
                    const i32 t = 2 * (12288 * sh.dd[2 + i] - 16384 * sh.dd[1 + i] + 4096 * sh.dd[i]);
                    const i32 fv = t >> 16, sq = 2 * fv * fv;
#pragma unroll
                    for (int j = 0; j < 4; j++)
                        if (j >= (32 * it) / 40 && j <= (32 * it + 31) / 40) e[j] += i / 40 == j ? sq : 0;
                }
                sh.red.a[l] = e[0]; sh.red.b[l] = e[1]; sh.red.c[l] = e[2]; sh.red.d[l] = e[3];
            LC3_LANES_END
            i32 be[4];
            be[0] = clamp64(warp_sum_wide(sh.red.a)); be[1] = clamp64(warp_sum_wide(sh.red.b));
            be[2] = clamp64(warp_sum_wide(sh.red.c)); be[3] = clamp64(warp_sum_wide(sh.red.d));
            as.filter_mem[0] = sh.dd[2 + 158];
            as.filter_mem[1] = sh.dd[2 + 159];
            i32 detected = as.position >= 2, pos = -1, acc = as.acc_energy;
#pragma unroll
            for (int j = 0; j < 4; j++) {
                const i32 lw = lo16(be[j] >> 1) & 0x7fff, hw = hi16(be[j]); // be / 8.5
                i32 q = l_mult0(lw, 30840) >> 15;
                q = l_mac0(q, hw, 30840) >> 2;
                if (q > acc) { pos = j; detected = 1; }
                acc = l_max(acc >> 2, be[j]);
            }
            as.acc_energy = acc; as.detected = detected; as.position = pos;
            LC3_SYNC; // dd is rewritten next frame
        }

        // ---- I. per-frame results, delay line for the next frame
        LC3_LANES(l)
            if (l < 24) sh.fir[l] = sh.s12[len + l];
            if (l == 0) {
                SV<i32> po = tile_view(b.pout, frame_tile(b, f, tile), PO_N, tl);
                po[PO_LTPF_BITS] = ltpf_bits; po[PO_LTPF_IDX0] = lt0; po[PO_LTPF_IDX1] = lt1; po[PO_LTPF_IDX2] = lt2;
                po[PO_ATTACK] = as.detected; po[PO_X_EXP] = x_exp;
            }
        LC3_LANES_END
        LC3_LANES(l)
            if (l < 24) sh.s12[l] = sh.fir[l];
        LC3_LANES_END
    }

    // ---- store the state
    LC3_LANES(l)
        if (l < 24) s16[S16_MEM_OUT + l] = sh.s12[l];
        if (l < 3) s16[S16_OLPA_S12 + l] = sh.m12[l];
        for (int k = l; k < olpa_len; k += 32) s16[S16_OLPA_BUF + k] = sh.olpa[ob][4 + k];
        for (int k = l; k < lt_len; k += 32) s16[S16_LTPF_BUF + k] = sh.ltpf[lb][4 + k];
        if (l == 0 && !PRE) { st[ST_MEM50_0] = m0; st[ST_MEM50_1] = m1; }
        if (l == 0) {
            st[ST_ATT_F0] = as.filter_mem[0]; st[ST_ATT_F1] = as.filter_mem[1]; st[ST_ATT_DET] = as.detected;
            st[ST_ATT_POS] = as.position; st[ST_ATT_SCAL] = as.scaling; st[ST_ATT_ACC] = as.acc_energy;
            st[ST_X_EXP] = x_exp; st[ST_RESAMP_EXP] = resamp_exp; st[ST_ATT_GEN] = att_gen;
            st[ST_OLPA_EXP] = olpa_exp; st[ST_OLPA_PITCH] = olpa_pitch;
            st[ST_LTPF_EXP] = ltpf_exp; st[ST_LTPF_NC] = ltpf_nc; st[ST_LTPF_MNC] = ltpf_mnc;
            st[ST_LTPF_ON] = ltpf_on; st[ST_LTPF_PITCH] = ltpf_pitch_mem;
        }
    LC3_LANES_END
}

// ---- resampler FIR as its own frame-parallel pass (16-bit input): one warp per (channel-stream, frame)
struct FirShared {
    alignas(4) i16 xs[64 + LC3K_MAX_N];  // xs[64 + i] = sample i of the frame
    alignas(4) i16 xs2[64 + LC3K_MAX_N]; // the same one sample earlier (xs2[k] = xs[k + 1]): an aligned pair for odd offsets
};

// Two taps as one word for the two-way dot product instruction (dp2a: two 16-bit samples times two BYTES): a tap is
// split into its unsigned low byte and its signed high byte, c = 256 * ch + cl; bytes 0,1 = cl of the two taps, 2,3 = ch.
LC3_HD u32 fir_pack_taps(i32 c0, i32 c1)
{
    return (u32)(c0 & 255) | ((u32)(c1 & 255) << 8) | ((u32)((c0 >> 8) & 255) << 16) | ((u32)((c1 >> 8) & 255) << 24);
}
LC3_HD u32 ld_u32(const i16 *p)
{
#if defined(__CUDA_ARCH__)
    return *reinterpret_cast<const u32 *>(p);
#else
    u32 v; memcpy(&v, p, 4); return v;
#endif
}
// acc + x.lo * (unsigned byte 0 of c) + x.hi * (unsigned byte 1 of c);  acc + x.lo * (signed byte 2) + x.hi * (signed byte 3)
LC3_HD i32 dp2a_lo_u8(u32 x, u32 c, i32 acc)
{
#if defined(__CUDA_ARCH__)
    i32 r; asm("dp2a.lo.s32.u32 %0, %1, %2, %3;" : "=r"(r) : "r"(x), "r"(c), "r"(acc)); return r;
#else
    return (i32)((u32)acc + (u32)((i32)(i16)(x & 0xffff) * (i32)(c & 255)) + (u32)(((i32)x >> 16) * (i32)((c >> 8) & 255)));
#endif
}
LC3_HD i32 dp2a_hi_s8(u32 x, u32 c, i32 acc)
{
#if defined(__CUDA_ARCH__)
    i32 r; asm("dp2a.hi.s32.s32 %0, %1, %2, %3;" : "=r"(r) : "r"(x), "r"(c), "r"(acc)); return r;
#else
    return (i32)((u32)acc + (u32)((i32)(i16)(x & 0xffff) * (i32)(int8_t)((c >> 16) & 255)) + (u32)(((i32)x >> 16) * ((i32)c >> 24)));
#endif
}

LC3_HD void fir_warp_body(const LC3K_Batch &b, int li, int f, FirShared &sh, const i16 *filt /* this rate's 240 taps */,
                          const u32 *filtp /* the same, fir_pack_taps of taps 2p, 2p + 1 */)
{
    if (li >= b.n_lanes) return;
    const int tile = li >> 5, tl = li & 31, N = b.N, rl = b.r12k8_mem_in_len, len = 32 * (b.dms / 25);
    const int upfac = rom_resamp_params[b.fs_idx * 4 + 0], delay2 = 2 * rom_resamp_params[b.fs_idx * 4 + 1];
    const int off_int = rom_resamp_params[b.fs_idx * 4 + 2], off_frac = rom_resamp_params[b.fs_idx * 4 + 3];
    RawPcm<i16> raw;
    raw.N = N;
    raw.hist = tile_view((const i16 *)b.pcm_t, (size_t)tile, (size_t)(b.n_frames + 1) * N, tl);
    raw.cur = (const i16 *)b.pcm + b.pcm_off[li] + (size_t)(b.frame0 + f) * N;
    raw.first = f == 0;
    LC3_LANES(l)
        int i = l - rl;
        for (; i < 0; i += 32) { const i16 v = (i16)raw(i); sh.xs[64 + i] = v; sh.xs2[63 + i] = v; } // previous frame: history or PCM
        for (; i < N; i += 32) { const i16 v = raw.cur[i]; sh.xs[64 + i] = v; sh.xs2[63 + i] = v; }
    LC3_LANES_END
    SV<i16> r = tile_view(b.r12, frame_tile(b, f, tile), 128, tl);
    // (see pitch_warp_body phase B for why plain arithmetic is exact.)  A lane owns outputs l, l + 32, l + 64, l + 96: the
    // up-sampling factor (2, 4 or 8) divides 32 * off_frac, so they share one polyphase branch and every coefficient is
    // fetched once for up to four products -- the pass is bound by shared-memory loads, not by the multiplies.
    LC3_LANES(l)
        const int nq = len >> 5; // 1, 2 or 4 outputs per lane
        const int t0 = l * off_frac, fr = t0 % upfac;
        const i16 *co = &filt[fr * delay2];
        const i16 *in0 = &sh.xs[64 + 1 + l * off_int + t0 / upfac - rl];
        const int step = 32 * off_int + (32 * off_frac) / upfac; // input distance between two outputs of the lane
        if (nq == 4) {
            // Two taps and two samples per instruction: the number of taps per branch, the branch's offset in the table and
            // `step` are even at every rate, so the pairs are aligned words (of xs or of its one-sample-shifted copy).
            // The low-byte and high-byte sums are exact (60 taps * 2^15 * 2^8 < 2^31) and 256 * hi + lo is the plain sum.
            const int e0 = (int)(in0 - sh.xs), sw = step >> 1, np = delay2 >> 1;
            const i16 *p0 = (e0 & 1) ? &sh.xs2[e0 - 1] : &sh.xs[e0];
            const u32 *cp = &filtp[(fr * delay2) >> 1];
            i32 l0 = 0, l1 = 0, l2 = 0, l3 = 0, h0 = 0, h1 = 0, h2 = 0, h3 = 0;
            for (int m = 0; m < np; m++) {
                const u32 c = cp[m];
                const u32 x0 = ld_u32(p0 + 2 * m), x1 = ld_u32(p0 + 2 * (m + sw)), x2 = ld_u32(p0 + 2 * (m + 2 * sw)), x3 = ld_u32(p0 + 2 * (m + 3 * sw));
                l0 = dp2a_lo_u8(x0, c, l0); h0 = dp2a_hi_s8(x0, c, h0);
                l1 = dp2a_lo_u8(x1, c, l1); h1 = dp2a_hi_s8(x1, c, h1);
                l2 = dp2a_lo_u8(x2, c, l2); h2 = dp2a_hi_s8(x2, c, h2);
                l3 = dp2a_lo_u8(x3, c, l3); h3 = dp2a_hi_s8(x3, c, h3);
            }
            const i32 a0 = (i32)(((u32)h0 << 8) + (u32)l0), a1 = (i32)(((u32)h1 << 8) + (u32)l1);
            const i32 a2 = (i32)(((u32)h2 << 8) + (u32)l2), a3 = (i32)(((u32)h3 << 8) + (u32)l3);
            r[l] = (i16)l_round(a0); r[l + 32] = (i16)l_round(a1); r[l + 64] = (i16)l_round(a2); r[l + 96] = (i16)l_round(a3);
        } else {
            for (int q = 0; q < nq; q++) {
                const i16 *in = in0 + q * step;
                i32 acc = 0;
                for (int m = 0; m < delay2; m++) acc += in[m] * co[m];
                r[l + 32 * q] = (i16)l_round(acc);
            }
        }
    LC3_LANES_END
}

// ---- the 50 Hz high-pass as its own pass (16-bit input): ONE LANE PER CHANNEL-STREAM walks the frames in order
// through the recursion (resamp12k8_fx.c:88-104), in place on b.r12.  It is the one part of the time-domain chain
// with no parallelism inside a stream; with a lane per stream it costs one warp instruction per 32 streams, and its run
// time is the length of the dependent chain per sample.  So (i) the recursion runs on plain wrapping additions -- five
// dependent instructions per sample instead of the saturating operators' thirteen -- with every addition's overflow
// condition collected off the chain; a frame in which any addition overflowed (none does on ordinary signals) is run
// again from its saved state with the saturating operators, so the result is the reference's in every case; (ii) the
// next frame's samples are fetched while the current frame's chain runs.
struct Hp50 { i32 n0, n1, n2, d0, d1; };

// one frame with the reference's operators
LC3_HD void hp50_frame_sat(const Hp50 &c, const i16 *in, i16 *out, int len, i32 &m0, i32 &m1)
{
    for (int n = 0; n < len; n++) {
        const i32 v = in[n];
        const i32 o = l_mac0(m0, c.n0, v);
        const i32 t = l_mac0(mpy32x16k(o, c.d0), c.n1, v);
        m0 = l_add(m1, l_shl_p(t, 1));
        m1 = l_mac0(mpy32x16k(o, c.d1), c.n2, v);
        out[n] = (i16)l_round(o);
    }
}

// HP50_BLOCK samples on wrapping additions; returns the accumulated overflow conditions in the sign bit
constexpr int HP50_BLOCK = 32; // long enough for the chain of one block to cover the latency of the next frame's loads
LC3_HD u32 hp50_run(const Hp50 &c, const i16 *in, i16 *out, i32 &m0_, i32 &m1_)
{
    u32 m0 = (u32)m0_, m1 = (u32)m1_, ovf = 0;
#pragma unroll
    for (int n = 0; n < HP50_BLOCK; n++) {
        const i32 v = in[n];
        const u32 p0 = (u32)(c.n0 * v), p1 = (u32)(c.n1 * v), p2 = (u32)(c.n2 * v);
        const u32 o = m0 + p0;
        ovf |= (m0 ^ o) & (p0 ^ o);
        const u32 x0 = (u32)mpy32x16k((i32)o, c.d0), t = x0 + p1;
        ovf |= (x0 ^ t) & (p1 ^ t);
        const u32 t2 = t << 1;
        ovf |= t ^ t2;
        const u32 a0 = m1 + t2;
        ovf |= (m1 ^ a0) & (t2 ^ a0);
        const u32 x1 = (u32)mpy32x16k((i32)o, c.d1), a1 = x1 + p2;
        ovf |= (x1 ^ a1) & (p2 ^ a1);
        m0 = a0; m1 = a1;
        out[n] = (i16)l_round((i32)o);
    }
    m0_ = (i32)m0; m1_ = (i32)m1;
    return ovf;
}

// Everything in registers: the walk goes block by block (HP50_BLOCK samples = one register tile), the NEXT block's samples are
// fetched before the current block's chain runs, results go straight back to b.r12, and the overflow check and the saturating
// re-run are per block (the block's inputs are still in registers, its starting state is saved).  No per-lane arrays in
// local memory: the earlier version kept a frame's samples and results there, and with one warp per scheduler nothing hid
// those round trips (0.55 -> see DESIGN.md section 4).
LC3_HD void scalar_body(const LC3K_Batch &b, int tile, int lane)
{
    const int li = tile * LC3K_LANES + lane;
    if (li >= b.n_lanes) return;
    SV<i32> st = slot_view(b.st32, b.slot[li], ST32_N);
    const int len = 32 * (b.dms / 25), nblk = len / HP50_BLOCK, total = b.n_frames * nblk; // blocks of the chunk
    i32 m0 = st[ST_MEM50_0], m1 = st[ST_MEM50_1];
    const Hp50 c = {rom_hp50_num[0], rom_hp50_num[1], rom_hp50_num[2], rom_hp50_den[0], rom_hp50_den[1]};
    i16 cur[HP50_BLOCK], nxt[HP50_BLOCK], res[HP50_BLOCK];
    auto row = [&](int blk) { return tile_view(b.r12, frame_tile(b, blk / nblk, tile), 128, lane) + (blk % nblk) * HP50_BLOCK; };
    {
        SV<i16> r = row(0);
#pragma unroll
        for (int j = 0; j < HP50_BLOCK; j++) cur[j] = r[j];
    }
    for (int blk = 0; blk < total; blk++) {
        SV<i16> r = row(blk), rn = row(blk + 1 < total ? blk + 1 : blk);
#pragma unroll
        for (int j = 0; j < HP50_BLOCK; j++) nxt[j] = rn[j]; // in flight during the chain below
        const i32 s0 = m0, s1 = m1;
        const u32 ovf = hp50_run(c, cur, res, m0, m1);
        if ((i32)ovf < 0) { m0 = s0; m1 = s1; hp50_frame_sat(c, cur, res, HP50_BLOCK, m0, m1); }
#pragma unroll
        for (int j = 0; j < HP50_BLOCK; j++) { r[j] = res[j]; cur[j] = nxt[j]; }
    }
    st[ST_MEM50_0] = m0; st[ST_MEM50_1] = m1;
}

} // namespace lc3k
