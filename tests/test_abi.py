"""The C-ABI library loads and exports every function include/lc3plus_b200.h declares; on a box without a
CUDA device the product fails loudly instead of falling back to a CPU path."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "lc3plus_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(lc3(?:plus)?_\w+)\s*\(", src)))


def test_header_symbols_exported():
    import audio_codec_b200 as ac
    L = ac.load_library()
    names = _declared()
    assert len(names) >= 30
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/lc3plus_b200.h but not exported"


def test_argument_checks_without_device():
    """Checks that need no device: NULL / alignment / samplerate / channels order (fx/lc3.c:126-129)."""
    import audio_codec_b200 as ac
    L = ac.load_library()
    buf = C.create_string_buffer(8192)
    base = (C.addressof(buf) + 15) & ~15
    assert L.lc3_enc_init(None, 48000, 1) == 3
    assert L.lc3_enc_init(C.c_void_p(base + 2), 48000, 1) == 11
    assert L.lc3_enc_init(C.c_void_p(base), 44000, 1) == 4
    assert L.lc3_enc_init(C.c_void_p(base), 48000, 3) == 5
    assert L.lc3_enc_get_size(48000, 2) > 0 and L.lc3_enc_get_size(48000, 2) <= 7064
    assert L.lc3_enc_get_size(12345, 1) == 0
    assert L.lc3_version() == (1 << 16) | (4 << 8) | 10
    assert L.lc3_samplerate_supported(44100) == 1 and L.lc3_samplerate_supported(11025) == 0
    assert L.lc3_enc_get_input_samples(None) == 0 and L.lc3_enc_get_num_bytes(None) == 0


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA device present")
    import audio_codec_b200 as ac
    with pytest.raises(ac.LC3Error) as e:
        ac.Encoder(48000, 1, 100, 128000)
    assert e.value.code == 1  # LC3_ERROR: no device, no fallback


def test_product_sources_do_not_reference_the_oracle():
    for d, _, files in os.walk(os.path.join(ROOT, "audio_codec_b200")):
        for f in files:
            if f.endswith((".c", ".h", ".cu", ".cuh", ".py")):
                txt = open(os.path.join(d, f)).read()
                assert "lc3o_" not in txt and "liblc3oracle" not in txt and "hostsim" not in txt.replace("tests/hostsim", ""), f
