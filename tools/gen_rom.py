#!/usr/bin/env python3
"""Generate the LC3plus encoder ROM (normative constants of ETSI TS 103 634) for this repo.

The table VALUES are data of the standard; they are loaded, not re-derived.  This script reads
them out of the compiled, unmodified reference (oracle/_ref/liblc3ref.so, built by
oracle/Makefile from /root/reference) through ctypes symbol lookup and re-packs them into our own
layout: per-(frame size, sample rate) tables are flattened into one array plus an offset table so
the device side never chases pointers.  It writes the same content twice:

    oracle/rom_tables.inc                      (included by the plain-C oracle)
    audio_codec_b200/csrc/rom_tables.inc       (included by the CUDA product code)

Every entry is `ROM_TAB(type, name, count) = { ... };`; each side defines ROM_TAB to the storage
class it wants.  Run once in the container that has /root/reference; the outputs are committed.

Reference symbols: fx/constants.c, fx/rom_basop_util.c (names in the comments below).
"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "oracle", "_ref", "liblc3ref.so")

CT = {"int8_t": C.c_int8, "int16_t": C.c_int16, "uint16_t": C.c_uint16,
      "int32_t": C.c_int32, "uint32_t": C.c_uint32}


def main():
    lib = C.CDLL(LIB)

    def arr(sym, ctype, n):
        return list((CT[ctype] * n).in_dll(lib, sym))

    def ptrs(sym, n):
        return list((C.c_void_p * n).in_dll(lib, sym))

    def deref(addr, ctype, n):
        return list((CT[ctype] * n).from_address(addr))

    out = []  # (ctype, name, values, comment)

    def emit(ctype, name, vals, comment=""):
        out.append((ctype, name, [int(v) for v in vals], comment))

    DMS = [("", 100), ("_5ms", 50), ("_2_5ms", 25)]

    # ---- MDCT windows (LowDelayShapes_n960*), flattened [dms][fs_idx]
    win_all, win_off, win_len, win_laz = [], [], [], []
    for sfx, _ in DMS:
        lens = arr("LowDelayShapes_n960_len" + sfx, "int16_t", 5)
        laz = arr("LowDelayShapes_n960_la_zeroes" + sfx, "int16_t", 5)
        pp = ptrs("LowDelayShapes_n960" + sfx, 5)
        for f in range(5):
            win_off.append(len(win_all))
            win_len.append(lens[f])
            win_laz.append(laz[f])
            win_all += deref(pp[f], "int16_t", lens[f])
    emit("int16_t", "rom_win", win_all, "LowDelayShapes_n960{,_5ms,_2_5ms}[fs_idx] flattened")
    emit("int16_t", "rom_win_off", win_off, "[dms_idx*5+fs_idx] offset into rom_win")
    emit("int16_t", "rom_win_len", win_len, "LowDelayShapes_n960_len*")
    emit("int16_t", "rom_win_laz", win_laz, "LowDelayShapes_n960_la_zeroes*")

    # ---- band tables
    nb = [[64] * 5, arr("bands_number_5ms", "int16_t", 5), arr("bands_number_2_5ms", "int16_t", 5)]
    bo_all, bo_off, bn_flat, one_max, two_max = [], [], [], [], []
    for d, (sfx, _) in enumerate(DMS):
        pp = ptrs("bands_offset" + sfx, 5)
        one = arr("bands_offset_with_one_max" + sfx, "int16_t", 5)
        two = arr("bands_offset_with_two_max" + sfx, "int16_t", 5)
        for f in range(5):
            bo_off.append(len(bo_all))
            bn_flat.append(nb[d][f])
            bo_all += deref(pp[f], "int16_t", nb[d][f] + 1)
            one_max.append(one[f])
            two_max.append(two[f])
    emit("int16_t", "rom_bands_offset", bo_all, "bands_offset{,_5ms,_2_5ms}[fs_idx] flattened")
    emit("int16_t", "rom_bands_offset_off", bo_off, "[dms_idx*5+fs_idx]")
    emit("int16_t", "rom_bands_number", bn_flat, "64 | bands_number_5ms | bands_number_2_5ms")
    emit("int16_t", "rom_bands_one_max", one_max, "bands_offset_with_one_max*")
    emit("int16_t", "rom_bands_two_max", two_max, "bands_offset_with_two_max*")
    emit("int16_t", "rom_bands_nrg_scale", arr("bands_nrg_scale", "int16_t", 32))
    emit("int16_t", "rom_inv_int", arr("InvIntTable", "int16_t", 32), "InvIntTable")

    # ---- bandwidth detector
    emit("int32_t", "rom_bw_thresh_quiet", arr("BW_thresh_quiet", "int32_t", 4))
    emit("int16_t", "rom_bw_thresh_quiet_exp", arr("BW_thresh_quiet_exp", "int16_t", 1))
    emit("int16_t", "rom_bw_thresh_brickwall", arr("BW_thresh_brickwall", "int16_t", 4))
    emit("int16_t", "rom_bw_cutoff_bin", arr("BW_cutoff_bin_all", "int16_t", 5))
    emit("int16_t", "rom_bw_cutoff_bits", arr("BW_cutoff_bits_all", "int16_t", 5))
    bd, ws, we = [], [], []
    for sfx, _ in DMS:
        bd += arr("BW_brickwall_dist" + sfx, "int16_t", 4)
        ps = ptrs("BW_warp_idx_start_all" + sfx, 4)
        pe = ptrs("BW_warp_idx_stop_all" + sfx, 4)
        for f in range(4):
            ws += deref(ps[f], "int16_t", 4)
            we += deref(pe[f], "int16_t", 4)
    emit("int16_t", "rom_bw_brickwall_dist", bd, "[dms_idx*4+k]")
    emit("int16_t", "rom_bw_warp_start", ws, "[(dms_idx*4+(fs_idx-1))*4+k]")
    emit("int16_t", "rom_bw_warp_stop", we, "[(dms_idx*4+(fs_idx-1))*4+k]")

    # ---- TNS
    ts, te = [], []
    for (sfx, _), n in zip(DMS, (6, 4, 2)):
        ps = ptrs("tns_subdiv_startfreq" + sfx, 5)
        pe = ptrs("tns_subdiv_stopfreq" + sfx, 5)
        for b in range(5):
            ts += deref(ps[b], "int16_t", n) + [0] * (6 - n)
            te += deref(pe[b], "int16_t", n) + [0] * (6 - n)
    emit("int16_t", "rom_tns_sub_start", ts, "tns_subdiv_startfreq*: [(dms_idx*5+bw)*6+k]")
    emit("int16_t", "rom_tns_sub_stop", te, "tns_subdiv_stopfreq*: [(dms_idx*5+bw)*6+k]")
    emit("int32_t", "rom_tns_acf_window", arr("tnsAcfWindow", "int32_t", 8))
    for nm in ("bits", "freq", "cumfreq"):
        emit("int16_t", "rom_tns_order_" + nm, arr("ac_tns_order_" + nm, "int16_t", 16), "[2][8]")
        emit("int16_t", "rom_tns_coef_" + nm, arr("ac_tns_coef_" + nm, "int16_t", 8 * 17), "[8][17]")
    emit("int16_t", "rom_tns_quant_pts", arr("tnsQuantPts", "int16_t", 17))
    emit("int16_t", "rom_tns_quant_thr", arr("tnsQuantThr", "int16_t", 16))

    # ---- spectrum arithmetic-coder model
    emit("int16_t", "rom_ari_esc_nb", arr("Tab_esc_nb", "int16_t", 4), "Tab_esc_nb")
    emit("int8_t", "rom_ari_lookup", arr("ari_spec_lookup", "int8_t", 4096))
    emit("uint16_t", "rom_ari_cumfreq", arr("ari_spec_cumfreq", "uint16_t", 64 * 17), "[64][17]")
    emit("uint16_t", "rom_ari_freq", arr("ari_spec_freq", "uint16_t", 64 * 17), "[64][17]")
    emit("uint16_t", "rom_ari_bits", arr("ari_spec_bits", "uint16_t", 64 * 17), "[64][17]")

    # ---- SNS
    pe, pee = ptrs("lpc_pre_emphasis", 5), ptrs("lpc_pre_emphasis_e", 5)
    a, b = [], []
    for f in range(5):
        a += deref(pe[f], "int16_t", 64)
        b += deref(pee[f], "int16_t", 64)
    emit("int16_t", "rom_sns_preemph", a, "lpc_pre_emphasis[fs_idx][64]")
    emit("int16_t", "rom_sns_preemph_e", b, "lpc_pre_emphasis_e[fs_idx][64]")
    gsz = arr("sns_gainSz", "int16_t", 4)
    gp = ptrs("sns_gaintabPtr", 4)
    g = []
    for j in range(4):
        g += deref(gp[j], "int16_t", gsz[j]) + [0] * (8 - gsz[j])
    emit("int16_t", "rom_sns_gain", g, "sns_gaintabPtr[j][i] padded to [4][8]")
    emit("int16_t", "rom_sns_gain_sz", gsz)
    emit("int16_t", "rom_sns_gain_msb_bits", arr("sns_gainMSBbits", "int16_t", 4))
    emit("int16_t", "rom_sns_gain_lsb_bits", arr("sns_gainLSBbits", "int16_t", 4))
    emit("int16_t", "rom_sns_kval", arr("sns_Kval", "int16_t", 8), "[4][2]")
    emit("uint32_t", "rom_sns_mpvq_sz", arr("sns_MPVQ_Sz", "uint32_t", 8), "[4][2]")
    emit("int16_t", "rom_sns_cb_lf", arr("st1SCF0_7_base5_32x8_Q14", "int16_t", 256))
    emit("int16_t", "rom_sns_cb_hf", arr("st1SCF8_15_base5_32x8_Q14", "int16_t", 256))
    emit("int16_t", "rom_isqrt_q16", arr("isqrt_Q16tab", "int16_t", 65))

    # ---- resampler / pitch
    rf = []
    pp = ptrs("resamp_filts", 5)
    for f in range(5):
        rf += deref(pp[f], "int16_t", 240)
    emit("int16_t", "rom_resamp_filt", rf, "resamp_filt_{8,16,24,32,48}k [fs_idx][240]")
    emit("int16_t", "rom_resamp_params", arr("resamp_params", "int16_t", 20), "[5][4]")
    emit("int16_t", "rom_hp50_num", arr("highpass50_filt_num", "int16_t", 3))
    emit("int16_t", "rom_hp50_den", arr("highpass50_filt_den", "int16_t", 2))
    emit("int16_t", "rom_olpa_weight", arr("olpa_ac_weighting", "int16_t", 98))
    emit("int16_t", "rom_ltpf_ac_interp", arr("ltpf_ac_interp_filt", "int16_t", 63), "[7][9]")
    emit("int16_t", "rom_ltpf_inter_filt", arr("inter_filter", "int16_t", 48), "inter_filter[0][4][12]")
    emit("int16_t", "rom_adjust_gain", arr("adjust_global_gain_tables", "int16_t", 25), "[5][5]")

    # ---- math helpers (rom_basop_util.c)
    emit("int16_t", "rom_ld_coeff", arr("ldCoeff", "int16_t", 7))
    emit("uint32_t", "rom_exp2_tab", arr("exp2_tab_long", "uint32_t", 32))
    emit("uint32_t", "rom_exp2w_tab", arr("exp2w_tab_long", "uint32_t", 32))
    emit("uint32_t", "rom_exp2x_tab", arr("exp2x_tab_long", "uint32_t", 32))
    emit("int32_t", "rom_isqrt16_tab", arr("ISqrtTable", "int32_t", 32))
    emit("int16_t", "rom_isqrt16_diff", arr("ISqrtDiffTable", "int16_t", 32))
    emit("int32_t", "rom_inv16_tab", arr("InvTable", "int32_t", 32))
    emit("int16_t", "rom_inv16_diff", arr("InvDiffTable", "int16_t", 32))
    emit("int32_t", "rom_isqrt32_tab", arr("isqrt_table", "int32_t", 130))
    emit("int32_t", "rom_log2_16_t1", arr("Log2_16_table1", "int32_t", 16))
    emit("int16_t", "rom_log2_16_t2", arr("Log2_16_table2", "int16_t", 16))
    emit("int32_t", "rom_invlog2_16_t1", arr("InvLog2_16_table1", "int32_t", 64))
    emit("int16_t", "rom_invlog2_16_t2", arr("InvLog2_16_table2", "int16_t", 64))

    # ---- DCT-IV / FFT twiddles.  PWord16 {re,im} pairs are stored interleaved.
    emit("int16_t", "rom_sin480", arr("SineTable480", "int16_t", 482), "SineTable480 re,im interleaved")
    emit("int16_t", "rom_sin320", arr("SineTable320", "int16_t", 322), "SineTable320 re,im interleaved")
    tw, tw_off = [], []
    for n in (20, 40, 60, 80, 120, 160, 240, 320, 480):
        tw_off.append(len(tw))
        tw += arr("SineWindow%d" % n, "int16_t", n)
    emit("int16_t", "rom_dct4_twiddle", tw, "SineWindow{20,40,60,80,120,160,240,320,480} interleaved, concatenated")
    emit("int16_t", "rom_dct4_twiddle_off", tw_off, "offset per N in the order above")
    emit("int16_t", "rom_rot320", arr("RotVector_320", "int16_t", 600))
    emit("int16_t", "rom_rot480", arr("RotVector_480", "int16_t", 900))
    emit("int32_t", "rom_rot40", arr("RotVector_40_32", "int32_t", 56))

    text = ["/* GENERATED by tools/gen_rom.py from the compiled ETSI reference -- do not edit.",
            " * Normative constants of ETSI TS 103 634 (LC3plus), re-packed into this repo's layout. */", ""]
    for ctype, name, vals, comment in out:
        if comment:
            text.append("/* %s */" % comment)
        text.append("ROM_TAB(%s, %s, %d) = {" % (ctype, name, len(vals)))
        line = "   "
        for v in vals:
            if ctype in ("uint32_t",):
                s = " 0x%08xu," % (v & 0xFFFFFFFF)
            elif ctype == "int32_t":
                s = " %d," % v if v != -2147483648 else " (-2147483647-1),"
            else:
                s = " %d," % v
            if len(line) + len(s) > 118:
                text.append(line)
                line = "   "
            line += s
        text.append(line)
        text.append("};")
        text.append("")
    body = "\n".join(text)
    decl = "\n".join(["/* GENERATED by tools/gen_rom.py -- declarations matching rom_tables.inc */"] +
                     ["ROM_DECL(%s, %s, %d);" % (c, n, len(v)) for c, n, v, _ in out]) + "\n"
    for rel in ("oracle/rom_tables_decl.inc", "audio_codec_b200/csrc/rom_tables_decl.inc"):
        with open(os.path.join(ROOT, rel), "w") as f:
            f.write(decl)
    for rel in ("oracle/rom_tables.inc", "audio_codec_b200/csrc/rom_tables.inc"):
        p = os.path.join(ROOT, rel)
        os.makedirs(os.path.dirname(p), exist_ok=True)
        with open(p, "w") as f:
            f.write(body)
        print("wrote", rel, len(body), "bytes")


if __name__ == "__main__":
    sys.exit(main())
