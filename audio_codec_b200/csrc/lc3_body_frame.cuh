// Frame-parallel passes: one lane encodes the spectral half of one channel-frame.
#pragma once
#include "lc3_fec.cuh"
#include "lc3_stages.cuh"
#include "lc3_mdct_tile.cuh"

namespace lc3k {

// ---- spectral front end: fx/enc_lc3.c:83-84, :119-175 (+ the state-free half of :181-185), everything in one
// lane-per-frame pass.  Used for the short frame lengths (N < 160); the longer ones run the tile MDCT kernel
// (lc3_mdct_tile.cuh: MDCT + band energies), then sns_body, then tns_body below.
template <int N, int BITS> LC3_HD void front_body(const LC3K_Batch &b, int f, int tile, int lane)
{
    const int li = tile * LC3K_LANES + lane;
    if (li >= b.n_lanes) return;
    const int slot = b.slot[li];
    SV<const i32> cfg = slot_view(b.cfg, slot, CFG_N);
    const size_t ft = frame_tile(b, f, tile);
    SV<const i32> po = tile_view((const i32 *)b.pout, ft, PO_N, lane);
    const i32 x_exp = po[PO_X_EXP], attack = po[PO_ATTACK];
    PcmView<BITS> xs = pcm_view<BITS>(b, tile, lane, f, x_exp);
    SV<i32> d = tile_view(b.spec, ft, N, lane);
    const i16 *bands = &rom_bands_offset[b.bands_off];

    SV<i32> fo = tile_view(b.fout, ft, FO_N, lane);
    i32 w[N];
    const i32 d_exp = mdct_frame<N>(xs, x_exp, &rom_win[b.W_off], b.W_len, b.mdct_mem_len, d, w);

    i32 en[64];
    const i32 en_exp = band_energy(en, d, d_exp, bands, b.fs_idx, b.nbands, b.dms_idx);
    i32 bw = b.fs_idx > 0 ? bw_detect(en, en_exp, b.fs_idx, b.dms_idx) : 0;

    i16 scf[16], scf_q[16], g[64], g_exp[64];
    i32 scf_idx[7];
    sns_compute(en, en_exp, b.fs_idx, b.nbands, scf, attack, 16384);
    sns_quantize(scf, scf_idx, scf_q);
    sns_interp(scf_q, g, g_exp, b.nbands);
    mdct_shape(d, g, g_exp, bands, b.nbands);

    const int cut = cfg[CFG_BW_CUTOFF_BIN];
    if (cut) { // lc3_enc_set_bandwidth was used (fx/cutoff_bandwidth.c:14-28, fx/enc_lc3.c:164-168)
        if (b.yLen > cut) {
            for (int i = -1; i < 3; i++) d[cut + i] = d[cut + i] >> (i + 2);
            for (int i = cut + 3; i < b.yLen; i++) d[i] = 0;
        }
        bw = imin(bw, cfg[CFG_BW_INDEX]);
    }

    TnsOut tns;
    if (b.dms_idx == 0) tns_coder<8>(tns, d, bw, cfg[CFG_LPC_WEIGHTING], b.nsub, b.dms_idx);
    else tns_coder<4>(tns, d, bw, cfg[CFG_LPC_WEIGHTING], b.nsub, b.dms_idx);

    SV<i32> en4 = tile_view(b.en4, ft, N / 4, lane);
    const i32 x_max = gain_energies(d, d_exp, b.yLen, en4);

    fo[FO_D_EXP] = d_exp; fo[FO_BW] = bw; fo[FO_TNS_BITS] = tns.bits; fo[FO_TNS_NFILT] = tns.nfilt;
    fo[FO_TNS_ORDER0] = tns.order[0]; fo[FO_TNS_ORDER1] = tns.order[1]; fo[FO_X_MAX] = x_max;
    for (int i = 0; i < 16; i++) fo[FO_TNS_IDX + i] = tns.idx[i];
    for (int i = 0; i < 7; i++) fo[FO_SCF + i] = scf_idx[i];
}

// ---- tile path, step 2: bandwidth detector and SNS from the band energies the tile MDCT kernel left in b.ben
// (fx/enc_lc3.c:127-160).  No spectrum access at all: the gains go back into b.ben, packed mantissa | exponent << 16.
LC3_HD void sns_body(const LC3K_Batch &b, int f, int tile, int lane)
{
    const int li = tile * LC3K_LANES + lane;
    if (li >= b.n_lanes) return;
    const size_t ft = frame_tile(b, f, tile);
    SV<i32> eo = tile_view(b.ben, ft, BEN_N, lane);
    SV<i32> fo = tile_view(b.fout, ft, FO_N, lane);
    const i32 attack = tile_view((const i32 *)b.pout, ft, PO_N, lane)[PO_ATTACK];
    i32 en[64];
    for (int i = 0; i < b.nbands; i++) en[i] = eo[i];
    const i32 en_exp = eo[BEN_EXP];
    const i32 bw = b.fs_idx > 0 ? bw_detect(en, en_exp, b.fs_idx, b.dms_idx) : 0;
    i16 scf[16], scf_q[16], g[64], g_exp[64];
    i32 scf_idx[7];
    sns_compute(en, en_exp, b.fs_idx, b.nbands, scf, attack, 16384);
    sns_quantize(scf, scf_idx, scf_q);
    sns_interp(scf_q, g, g_exp, b.nbands);
    for (int i = 0; i < b.nbands; i++) eo[i] = (i32)((u32)g[i] & 0xffffu) | (i32)((u32)(i32)g_exp[i] << 16);
    fo[FO_BW] = bw;
    for (int i = 0; i < 7; i++) fo[FO_SCF + i] = scf_idx[i];
}

// ---- tile path, step 3: spectral shaping, bandwidth cut-off, TNS, per-quadruple log-energies
// (fx/enc_lc3.c:162-175, first half of :181-185)
template <int N> LC3_HD void tns_body(const LC3K_Batch &b, int f, int tile, int lane)
{
    const int li = tile * LC3K_LANES + lane;
    if (li >= b.n_lanes) return;
    SV<const i32> cfg = slot_view(b.cfg, b.slot[li], CFG_N);
    const size_t ft = frame_tile(b, f, tile);
    SV<const i32> eo = tile_view((const i32 *)b.ben, ft, BEN_N, lane);
    SV<i32> fo = tile_view(b.fout, ft, FO_N, lane);
    SV<i32> d = tile_view(b.spec, ft, N, lane);
    const i16 *bands = &rom_bands_offset[b.bands_off];
    const i32 d_exp = fo[FO_D_EXP];
    i32 bw = fo[FO_BW];
    i16 g[64], g_exp[64];
    for (int i = 0; i < b.nbands; i++) { const i32 v = eo[i]; g[i] = (i16)(v & 0xffff); g_exp[i] = (i16)(v >> 16); }
    mdct_shape(d, g, g_exp, bands, b.nbands);

    const int cut = cfg[CFG_BW_CUTOFF_BIN];
    if (cut) { // lc3_enc_set_bandwidth was used (fx/cutoff_bandwidth.c:14-28, fx/enc_lc3.c:164-168)
        if (b.yLen > cut) {
            for (int i = -1; i < 3; i++) d[cut + i] = d[cut + i] >> (i + 2);
            for (int i = cut + 3; i < b.yLen; i++) d[i] = 0;
        }
        bw = imin(bw, cfg[CFG_BW_INDEX]);
    }

    TnsOut tns;
    if (b.dms_idx == 0) tns_coder<8>(tns, d, bw, cfg[CFG_LPC_WEIGHTING], b.nsub, b.dms_idx);
    else tns_coder<4>(tns, d, bw, cfg[CFG_LPC_WEIGHTING], b.nsub, b.dms_idx);

    SV<i32> en4 = tile_view(b.en4, ft, N / 4, lane);
    const i32 x_max = gain_energies(d, d_exp, b.yLen, en4);

    fo[FO_BW] = bw; fo[FO_TNS_BITS] = tns.bits; fo[FO_TNS_NFILT] = tns.nfilt;
    fo[FO_TNS_ORDER0] = tns.order[0]; fo[FO_TNS_ORDER1] = tns.order[1]; fo[FO_X_MAX] = x_max;
    for (int i = 0; i < 16; i++) fo[FO_TNS_IDX + i] = tns.idx[i];
}

// ---- back end: fx/enc_lc3.c:196-242 (adjust gain, final quantisation, residual, noise factor, entropy coding)
// The finished frame is left in `bytes` (see lc3_bits.cuh); returns its size and, in dst, where the caller must copy it
// (0: lane not in use).
template <class B>
LC3_HD i32 back_body(const LC3K_Batch &b, int f, int tile, int lane, const AriTabs &T, B bytes, u8 *&dst)
{
    const int li = tile * LC3K_LANES + lane;
    dst = nullptr;
    if (li >= b.n_lanes) return 0;
    const int slot = b.slot[li];
    SV<const i32> cfg = slot_view(b.cfg, slot, CFG_N);
    const size_t ft = frame_tile(b, f, tile);
    SV<const i32> po = tile_view((const i32 *)b.pout, ft, PO_N, lane);
    SV<const i32> fo = tile_view((const i32 *)b.fout, ft, FO_N, lane);
    SV<const i32> ro = tile_view((const i32 *)b.rout, ft, RO_N, lane);
    SV<const i32> d = tile_view((const i32 *)b.spec, ft, b.N, lane);
    const i32 target_bytes = cfg[CFG_TARGET_BYTES], total_bits = cfg[CFG_TOTAL_BITS], qgain_off = cfg[CFG_QGAIN_OFF];
    const i32 lpc_weighting = cfg[CFG_LPC_WEIGHTING];
    const i32 d_exp = fo[FO_D_EXP], bw = fo[FO_BW], tbits_quant = ro[RO_TBITS_QUANT];

    // gain index after adjustment, quantised bins and their summary come from the tile quantiser (lc3_quant_tile.cuh)
    const i32 qgain = ro[RO_QGAIN];
    i32 gain, gain_e;
    gain_from_index(s_add(qgain, qgain_off), gain, gain_e);
    SV<const i16> xq = tile_view((const i16 *)b.xq, ft, b.N, lane);
    QuantOut q;
    q.lastnz = ro[RO_LASTNZ]; q.lsbMode = ro[RO_LSBMODE]; q.nBits2 = ro[RO_NBITS2]; q.nBits = 0;

    u32 res[LC3K_MAX_N / 32 + 1], lsbs[2 * LC3K_MAX_N / 32 + 1];
    i32 nres = 0;
    if (q.lsbMode == 0) nres = residual_bits(d_exp, d, xq, gain, gain_e, b.yLen, tbits_quant, q.nBits2, res);
    const i32 fac_ns = noise_factor(d_exp, d, xq, gain, gain_e, bw, b.dms, target_bytes);

    SideInfo si;
    si.bw = bw; si.bw_bits = b.bw_bits; si.lastnz = q.lastnz; si.lsbMode = q.lsbMode; si.gg = qgain;
    TnsOut tns;
    tns.nfilt = si.nfilt = fo[FO_TNS_NFILT];
    tns.order[0] = si.tns_order[0] = fo[FO_TNS_ORDER0];
    tns.order[1] = si.tns_order[1] = fo[FO_TNS_ORDER1];
    tns.bits = fo[FO_TNS_BITS];
    for (int i = 0; i < 16; i++) tns.idx[i] = (i16)fo[FO_TNS_IDX + i];
    si.ltpf_idx[0] = po[PO_LTPF_IDX0]; si.ltpf_idx[1] = po[PO_LTPF_IDX1]; si.ltpf_idx[2] = po[PO_LTPF_IDX2];
    si.fac_ns = fac_ns;
    si.bfi_ext = b.bfi ? b.bfi[(size_t)(b.frame0 + f) * b.n_lanes + li] : 0;
    for (int i = 0; i < 7; i++) si.scf[i] = fo[FO_SCF + i];

    bytes.zero(target_bytes);
    BackWriter<B> w;
    side_info(w, bytes, total_bits, b.yLen, si, b.fs_idx);
    const i32 rateFlag = s_sub(total_bits, s_add(160, s_imult(b.fs_idx, 160))) > 0 ? 2 << 8 : 0;
    // partial concealment only with error protection per channel (fx/enc_lc3.c:247-259)
    const i32 n_pc = (cfg[CFG_EP] != 0 && (cfg[CFG_EP] >> 8) == 0 && cfg[CFG_N_PCCW] > 0) ? cfg[CFG_N_PC] : 0;
#if defined(__CUDA_ARCH__)
    const bool any_pc = __any_sync(__activemask(), n_pc > 0);
#else
    const bool any_pc = n_pc > 0;
#endif
    const i32 b_left = any_pc ? ari_encode<true>(bytes, w, total_bits, xq, b.yLen, rateFlag, tns, q.lastnz, res, nres, q.lsbMode,
                                                 lpc_weighting, T, lsbs, n_pc)
                              : ari_encode<false>(bytes, w, total_bits, xq, b.yLen, rateFlag, tns, q.lastnz, res, nres, q.lsbMode,
                                                  lpc_weighting, T, lsbs, 0);
    if (n_pc > 0) reorder_payload(bytes, n_pc, b_left);

    dst = b.out + b.out_off[li] + (size_t)(b.frame0 + f) * b.out_stride[li];
    return target_bytes;
}

// ---- error protection of the lane's finished payload, in place in the caller's output (fx/enc_lc3.c:282-312)
LC3_HD void fec_body(const LC3K_Batch &b, int f, int tile, int lane, const FecTabs &T)
{
    const int li = tile * LC3K_LANES + lane;
    if (li >= b.n_lanes) return;
    SV<const i32> cfg = slot_view(b.cfg, b.slot[li], CFG_N);
    const i32 ep = cfg[CFG_EP], slot_bytes = cfg[CFG_SLOT_BYTES];
    if (ep == 0 || slot_bytes == 0) return; // off, or channel 1 of a jointly coded frame
    u8 *io = b.out + b.out_off[li] + (size_t)(b.frame0 + f) * b.out_stride[li];
    fec_encode_slot(T, ep & 15, (ep >> 4) & 15, io, cfg[CFG_EP_DATA_BYTES], slot_bytes, cfg[CFG_N_PCCW]);
}

} // namespace lc3k
