// Per-stream passes: the two places where frame n+1 needs frame n's result.
#pragma once
#include "lc3_stages.cuh"

namespace lc3k {

// ---- time-domain pass: fx/enc_lc3.c:65-114 (scale_signal24, resamp12k8, olpa, ltpf_coder, attack_detector)
template <int BITS> LC3_HD void pitch_body(const LC3K_Batch &b, int tile, int lane)
{
    const int li = tile * LC3K_LANES + lane;
    if (li >= b.n_lanes) return;
    const int slot = b.slot[li];
    SV<i32> st = slot_view(b.st32, slot, ST32_N);
    SV<i16> s16 = slot_view(b.st16, slot, ST16_N);
    SV<const i32> cfg = slot_view(b.cfg, slot, CFG_N);
    const i32 ltpf_enable = cfg[CFG_LTPF_ENABLE], attack_handling = cfg[CFG_ATTACK];

    PitchState ps;
    AttackState as;
    ps.r12k8_mem_50[0] = st[ST_MEM50_0]; ps.r12k8_mem_50[1] = st[ST_MEM50_1];
    for (int k = 0; k < 24; k++) ps.r12k8_mem_out[k] = s16[S16_MEM_OUT + k];
    for (int k = 0; k < 3; k++) ps.olpa_mem_s12k8[k] = s16[S16_OLPA_S12 + k];
    const int olpa_len = b.dms == 25 ? 130 : 114;
    for (int k = 0; k < olpa_len; k++) ps.olpa_buf[k] = s16[S16_OLPA_BUF + k];
    for (int k = 0; k < b.ltpf_mem_in_len; k++) ps.ltpf_buf[k] = s16[S16_LTPF_BUF + k];
    ps.olpa_exp = st[ST_OLPA_EXP]; ps.olpa_pitch = st[ST_OLPA_PITCH];
    ps.ltpf_exp = st[ST_LTPF_EXP]; ps.ltpf_normcorr = st[ST_LTPF_NC]; ps.ltpf_mem_normcorr = st[ST_LTPF_MNC];
    ps.ltpf_on = st[ST_LTPF_ON]; ps.ltpf_pitch = st[ST_LTPF_PITCH];
    as.filter_mem[0] = st[ST_ATT_F0]; as.filter_mem[1] = st[ST_ATT_F1]; as.detected = st[ST_ATT_DET];
    as.position = st[ST_ATT_POS]; as.scaling = st[ST_ATT_SCAL]; as.acc_energy = st[ST_ATT_ACC];
    if (!attack_handling) { // the reference clears the detector whenever set_bitrate switches it off (setup_enc_lc3.c:252-275)
        as.filter_mem[0] = as.filter_mem[1] = 0; as.detected = as.position = as.scaling = as.acc_energy = 0;
    }
    i32 x_exp = st[ST_X_EXP], resamp_exp = st[ST_RESAMP_EXP];

    const int len = 32 * (b.dms / 25);
    for (int f = 0; f < b.n_frames; f++) {
        if (BITS == 24) { // block exponent of this frame (fx/scale_signal24_fx.c:13-92)
            PcmView<24> raw = pcm_view<24>(b, tile, lane, f, 0);
            Range r0, r1, r2;
            range_init(r0); range_init(r1); range_init(r2);
            for (int i = 0; i < b.N; i++) range_add(r0, raw.p[i]);
            for (int i = -b.mdct_mem_len; i < 0; i++) range_add(r1, raw.p[i]);
            for (int i = -b.r12k8_mem_in_len; i < 0; i++) range_add(r2, raw.p[i]);
            Range r3, r4, r5;
            range_init(r3); range_init(r4); range_init(r5);
            range_add(r3, ps.r12k8_mem_50[0]); range_add(r3, ps.r12k8_mem_50[1]);
            for (int k = 0; k < 24; k++) range_add(r4, ps.r12k8_mem_out[k]);
            for (int k = 0; k < 3; k++) range_add(r5, ps.olpa_mem_s12k8[k]);
            i32 e = 7;
            e = s_max(e, s_sub(23, range_headroom32(r0, 32)));
            e = s_max(e, s_sub(23, range_headroom32(r1, 32)));
            e = s_max(e, s_sub(23, range_headroom32(r2, 32)));
            e = s_max(e, s_sub(s_sub(resamp_exp, 2), range_headroom32(r3, 32)));
            e = s_max(e, s_sub(s_sub(resamp_exp, 2), range_headroom16(r4, 16)));
            e = s_max(e, s_sub(s_sub(resamp_exp, 2), range_headroom16(r5, 16)));
            x_exp = e;
            const i32 sh = s_sub(s_sub(resamp_exp, 2), x_exp);
            resamp_exp = s_add(x_exp, 2);
            if (sh) {
                ps.r12k8_mem_50[0] = l_shl(ps.r12k8_mem_50[0], sh);
                ps.r12k8_mem_50[1] = l_shl(ps.r12k8_mem_50[1], sh);
                for (int k = 0; k < 24; k++) ps.r12k8_mem_out[k] = (i16)s_shl(ps.r12k8_mem_out[k], sh);
                for (int k = 0; k < 3; k++) ps.olpa_mem_s12k8[k] = (i16)s_shl(ps.olpa_mem_s12k8[k], sh);
            }
        }
        PcmView<BITS> xs = pcm_view<BITS>(b, tile, lane, f, x_exp);

        i16 s12[24 + 128];
        resamp12k8(xs, b.r12k8_mem_in_len, ps, s12, len, b.fs_idx);
        i32 normcorr;
        const i32 pitch = olpa(ps, s12, len, &normcorr, resamp_exp, b.dms);
        i32 lt[3];
        const i32 ltpf_bits = ltpf_coder(ps, pitch, ltpf_enable, b.ltpf_mem_in_len, lt, s12, len, normcorr, resamp_exp, b.dms);
        if (attack_handling) attack_detector(as, xs, b.N, b.fs, s_sub(x_exp, 15));

        SV<i32> po = tile_view(b.pout, frame_tile(b, f, tile), PO_N, lane);
        po[PO_LTPF_BITS] = ltpf_bits; po[PO_LTPF_IDX0] = lt[0]; po[PO_LTPF_IDX1] = lt[1]; po[PO_LTPF_IDX2] = lt[2];
        po[PO_ATTACK] = as.detected; po[PO_X_EXP] = x_exp;
    }

    st[ST_MEM50_0] = ps.r12k8_mem_50[0]; st[ST_MEM50_1] = ps.r12k8_mem_50[1];
    for (int k = 0; k < 24; k++) s16[S16_MEM_OUT + k] = ps.r12k8_mem_out[k];
    for (int k = 0; k < 3; k++) s16[S16_OLPA_S12 + k] = ps.olpa_mem_s12k8[k];
    for (int k = 0; k < olpa_len; k++) s16[S16_OLPA_BUF + k] = ps.olpa_buf[k];
    for (int k = 0; k < b.ltpf_mem_in_len; k++) s16[S16_LTPF_BUF + k] = ps.ltpf_buf[k];
    st[ST_OLPA_EXP] = ps.olpa_exp; st[ST_OLPA_PITCH] = ps.olpa_pitch;
    st[ST_LTPF_EXP] = ps.ltpf_exp; st[ST_LTPF_NC] = ps.ltpf_normcorr; st[ST_LTPF_MNC] = ps.ltpf_mem_normcorr;
    st[ST_LTPF_ON] = ps.ltpf_on; st[ST_LTPF_PITCH] = ps.ltpf_pitch;
    st[ST_ATT_F0] = as.filter_mem[0]; st[ST_ATT_F1] = as.filter_mem[1]; st[ST_ATT_DET] = as.detected;
    st[ST_ATT_POS] = as.position; st[ST_ATT_SCAL] = as.scaling; st[ST_ATT_ACC] = as.acc_energy;
    st[ST_X_EXP] = x_exp; st[ST_RESAMP_EXP] = resamp_exp;
}

// ---- rate-loop pass: fx/enc_lc3.c:180-192 (estimate global gain + first quantisation pass bit count)
LC3_HD void rate_body(const LC3K_Batch &b, int tile, int lane, const u8 *ari_lut, const uint16_t *ari_bitstab)
{
    const int li = tile * LC3K_LANES + lane;
    if (li >= b.n_lanes) return;
    const int slot = b.slot[li];
    SV<i32> st = slot_view(b.st32, slot, ST32_N);
    SV<const i32> cfg = slot_view(b.cfg, slot, CFG_N);
    const i32 tbits_init = cfg[CFG_TBITS_INIT], total_bits = cfg[CFG_TOTAL_BITS], qgain_off = cfg[CFG_QGAIN_OFF];
    RateState rs = {st[ST_TBITS_OFF], st[ST_MEM_TBITS], st[ST_MEM_SPECBITS]};

    i16 xq[LC3K_MAX_N];
    for (int f = 0; f < b.n_frames; f++) {
        const size_t ft = frame_tile(b, f, tile);
        SV<const i32> fo = tile_view((const i32 *)b.fout, ft, FO_N, lane);
        SV<const i32> po = tile_view((const i32 *)b.pout, ft, PO_N, lane);
        SV<const i32> d = tile_view((const i32 *)b.spec, ft, b.N, lane);
        SV<const i32> en = tile_view((const i32 *)b.en4, ft, b.N / 4, lane);
        const i32 d_exp = fo[FO_D_EXP];
        const i32 tbits_quant = s_sub(tbits_init, s_add(fo[FO_TNS_BITS], po[PO_LTPF_BITS]));
        i32 gain, gain_e, qgain, qgain_min;
        est_gain(en, b.yLen >> 2, fo[FO_X_MAX], d_exp, tbits_quant, qgain_off, rs, gain, gain_e, qgain, qgain_min);
        QuantOut q;
        quantize<false>(q, d, d_exp, gain, gain_e, xq, b.yLen, tbits_quant, total_bits, b.fs_idx, (i16 *)0, -1,
                        ari_lut, ari_bitstab);
        rs.mem_specbits = q.nBits;
        SV<i32> ro = tile_view(b.rout, ft, RO_N, lane);
        ro[RO_QGAIN] = qgain; ro[RO_QGAIN_MIN] = qgain_min; ro[RO_NBITS1] = q.nBits; ro[RO_TBITS_QUANT] = tbits_quant;
    }
    st[ST_TBITS_OFF] = rs.tbits_off; st[ST_MEM_TBITS] = rs.mem_tbits; st[ST_MEM_SPECBITS] = rs.mem_specbits;
}

} // namespace lc3k
