"""Range facts about the ROM that the kernels rely on where they drop a saturation the reference carries."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _tab(name):
    s = open(os.path.join(ROOT, "audio_codec_b200", "csrc", "rom_tables.inc")).read()
    m = re.search(r"ROM_TAB\(\w+, " + name + r", (\d+)\) = \{(.*?)\};", s, re.S)
    v = [int(x, 0) for x in re.findall(r"-?(?:0x[0-9a-fA-F]+|\d+)", m.group(2))]
    assert len(v) == int(m.group(1))
    return v


def test_resampler_taps_cannot_overflow_32_bits():
    """lc3_pitch_warp.cuh phase B: sum |taps| of every polyphase branch < 2^16, so |sum x*c| < 2^31 for |x| <= 2^15."""
    f, p = _tab("rom_resamp_filt"), _tab("rom_resamp_params")
    for fs in range(5):
        up, delay = p[fs * 4], p[fs * 4 + 1]
        for fr in range(up):
            co = f[fs * 240 + fr * delay * 2: fs * 240 + (fr + 1) * delay * 2]
            assert sum(abs(c) for c in co) * 32768 < 2 ** 31


def test_ltpf_interpolation_taps_cannot_overflow_on_block_scaled_input():
    """lc3_pitch_warp.cuh phase G: the 3-/4-tap filtered current and predicted frames are computed with plain arithmetic.
    The LTPF buffers are block-scaled to three bits of headroom (|x| <= 2^12), so 2 * sum |taps| * 2^12 + 2^15 < 2^31."""
    f = _tab("rom_ltpf_inter_filt")
    for ph in range(4):
        assert 2 * sum(abs(c) for c in f[ph * 12:(ph + 1) * 12]) * 4096 + 0x8000 < 2 ** 31


def test_escape_rows_are_multiples_of_1024():
    """lc3_bits.cuh ari_encode and lc3_rate_warp.cuh write the escape row offset as min(q, 3) << 10."""
    assert list(_tab("rom_ari_esc_nb")) == [0, 1024, 2048, 3072]


def test_product_and_oracle_rom_identical():
    a = open(os.path.join(ROOT, "audio_codec_b200", "csrc", "rom_tables.inc")).read()
    b = open(os.path.join(ROOT, "oracle", "rom_tables.inc")).read()
    assert a == b


def test_constants_used_with_unchecked_multiply_are_not_minus_one():
    """mpy32x16k / mpy32x32k skip the (-1 x -1) saturation test: their constant operands must differ from the minimum."""
    for name in ("rom_hp50_num", "rom_hp50_den", "rom_ltpf_inter_filt", "rom_tns_quant_pts", "rom_win"):
        assert -32768 not in _tab(name), name
    # decoder-only tables (tools/gen_rom_dec.py): the LTPF synthesis taps multiply without L_mult's -32768 x -32768 case
    import re
    txt = open(os.path.join(ROOT, "audio_codec_b200", "csrc", "rom_dec_tables.inc")).read()
    for name in ("rom_ltpf_dec_inter", "rom_ltpf_dec_tilt"):
        body = re.search(r"ROM_TAB\(\w+, %s, \d+\) = \{(.*?)\};" % name, txt, re.S).group(1)
        assert -32768 not in [int(v) for v in re.findall(r"-?\d+", body)], name
    assert txt == open(os.path.join(ROOT, "oracle", "rom_dec_tables.inc")).read()


def test_tns_subdivisions_tile_the_filter_bins():
    """lc3_tns.cuh takes max |x| over a TNS filter's bins from the range scans of its subdivisions: for every frame
    duration and bandwidth they must cover [start frequency, stop) without gaps (fx/tns_coder_fx.c:66-75, :288)."""
    lo_t, hi_t, cut = _tab("rom_tns_sub_start"), _tab("rom_tns_sub_stop"), _tab("rom_bw_cutoff_bin")
    for dms_idx in range(3):
        nsub = 3 if dms_idx == 0 else 2
        for bw in range(5):
            xlen = cut[bw] >> dms_idx
            nfilt = 2 if (bw >= 3 and dms_idx <= 1) else 1
            start = [12 >> dms_idx, xlen >> 1]
            base = (dms_idx * 5 + bw) * 6
            for j in range(nfilt):
                stop = start[1] if (nfilt == 2 and j == 0) else xlen
                lo = [lo_t[base + nsub * j + sd] for sd in range(nsub)]
                hi = [hi_t[base + nsub * j + sd] for sd in range(nsub)]
                assert lo[0] <= start[j] and hi[-1] >= stop, (dms_idx, bw, j, lo, hi, start[j], stop)
                assert all(hi[sd] >= lo[sd + 1] for sd in range(nsub - 1)), (dms_idx, bw, j, lo, hi)


def test_cumulative_frequency_rows_rise_monotonically_from_zero():
    """the decoder finds a symbol as the largest r with low >= help * cumfreq[r] in two rounds of comparisons instead of the
    reference's four-step bisection (lc3_dec_frame.cuh rdc_decode17): the two agree because every row is non-decreasing from 0"""
    for name, rows in (("rom_ari_cumfreq", 64), ("rom_tns_coef_cumfreq", 8)):
        v = _tab(name)
        assert len(v) == rows * 17
        for r in range(rows):
            row = v[r * 17:(r + 1) * 17]
            assert row[0] == 0 and all(a <= b for a, b in zip(row, row[1:])), (name, r)
