"""Kernel bodies compiled for the host (tests/hostsim, test tooling) against the oracle.  This is how the
stage code of audio_codec_b200/csrc is debugged where there is no GPU; the GPU tests repeat the comparison
through the real kernels and the C ABI."""
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "hostsim"))
import run as hs  # noqa: E402
from audio_codec_b200.synth import make_streams, make_torture  # noqa: E402

CASES = [(48000, 100, 128000, 16), (16000, 50, 32000, 16), (48000, 25, 128000, 16), (8000, 100, 16000, 16),
         (24000, 50, 48000, 24), (32000, 100, 80000, 16), (44100, 100, 96000, 16), (8000, 25, 64000, 16),
         (48000, 100, 128000, 24), (32000, 25, 96000, 16), (24000, 100, 16000, 16), (48000, 50, 48000, 16)]


@pytest.fixture(scope="module")
def L():
    return hs.lib()


@pytest.mark.parametrize("fs,dms,br,bits", CASES)
def test_bodies_equal_oracle(L, fs, dms, br, bits):
    N = (48000 if fs == 44100 else fs) // 100 * dms // 100
    nf, S = 16, 9
    pcm = make_streams(S, nf * N, fs, N, bits, silence_frames=6)
    a = hs.sim_encode(L, fs, dms, [(br, nf)], pcm, bits)[0]
    b = hs.oracle_encode(fs, dms, [(br, nf)], pcm, bits)[0]
    assert np.array_equal(a, b)


def test_bodies_bitrate_switching_and_lost_frames(L):
    fs, dms = 48000, 100
    sched = [(128000, 5), (32000, 5), (320000, 5), (96000, 5)]
    pcm = make_streams(3, 20 * 480, fs, 480, 16, first_stream=2)
    bfi = [0, 0, 1, 0, 0] * 4
    a = hs.sim_encode(L, fs, dms, sched, pcm, 16, bfi)
    b = hs.oracle_encode(fs, dms, sched, pcm, 16, bfi)
    for x, y in zip(a, b):
        assert np.array_equal(x, y)


def test_host_config_matches_reference_getters(L):
    """lc3_config.c (host half of the product) against the compiled reference, every fs x dms x a bitrate grid."""
    import ctypes as C
    import lc3ref
    if not lc3ref.have_ref():
        pytest.skip("reference library absent")
    R = lc3ref.ref_lib()

    class Cfg(C.Structure):
        _fields_ = [("w", C.c_int32 * 64)]
    L.lc3cfg_init.argtypes = [C.c_void_p, C.c_int, C.c_int]
    L.lc3cfg_set_frame_dms.argtypes = [C.c_void_p, C.c_int]
    L.lc3cfg_set_bitrate.argtypes = [C.c_void_p, C.c_int]
    L.lc3cfg_num_bytes.argtypes = [C.c_void_p]
    L.lc3cfg_real_bitrate.argtypes = [C.c_void_p]
    for fs in (8000, 16000, 24000, 32000, 44100, 48000):
        for ch in (1, 2):
            for dms in (25, 50, 100):
                for br in list(range(8000, 700000, 23000)) + [16000, 32000, 64000, 128000, 256000, 320000]:
                    ref = lc3ref.RefEnc(fs, ch, dms)
                    c = Cfg()
                    assert L.lc3cfg_init(C.byref(c), fs, ch) == 0
                    assert L.lc3cfg_set_frame_dms(C.byref(c), dms) == 0
                    r0, r1 = ref.set_bitrate(br), L.lc3cfg_set_bitrate(C.byref(c), br)
                    assert r0 == r1, (fs, ch, dms, br)
                    if r0 == 0:
                        assert ref.num_bytes == L.lc3cfg_num_bytes(C.byref(c))
                        assert R.lc3_enc_get_real_bitrate(ref.h) == L.lc3cfg_real_bitrate(C.byref(c))


@pytest.mark.parametrize("fs,dms,br,bits", [(48000, 100, 128000, 16), (48000, 100, 320000, 24), (16000, 50, 32000, 16),
                                            (48000, 25, 256000, 16), (32000, 100, 64000, 16), (8000, 100, 16000, 24),
                                            (24000, 25, 96000, 24), (44100, 100, 160000, 16)])
def test_bodies_equal_oracle_on_adversarial_signals(L, fs, dms, br, bits):
    """Rail-to-rail noise, squares, DC, impulses, steps: the inputs that reach the saturating operators."""
    N = (48000 if fs == 44100 else fs) // 100 * dms // 100
    nf = 14
    pcm = make_torture(nf * N, fs, bits)
    a = hs.sim_encode(L, fs, dms, [(br, nf)], pcm, bits)[0]
    b = hs.oracle_encode(fs, dms, [(br, nf)], pcm, bits)[0]
    bad = np.argwhere((a != b).any(2))
    assert len(bad) == 0, f"first mismatch (stream, frame) = {bad[0]}"


@pytest.mark.parametrize("fs,dms,br,bits", [(48000, 100, 128000, 16), (32000, 100, 32000, 16), (48000, 50, 96000, 24), (16000, 25, 64000, 16)])
def test_bodies_equal_oracle_on_transient_rich_signals(L, fs, dms, br, bits):
    """Gated tones, noise bursts and clicks: temporal noise shaping, the attack detector (low rates) and pitch
    changes are active in most frames, unlike in the reference's stationary stimulus."""
    from audio_codec_b200.synth import make_bursts
    N = fs // 100 * dms // 100
    nf = 30
    pcm = make_bursts(12, nf * N, fs, bits)
    a = hs.sim_encode(L, fs, dms, [(br, nf)], pcm, bits)[0]
    b = hs.oracle_encode(fs, dms, [(br, nf)], pcm, bits)[0]
    bad = np.argwhere((a != b).any(2))
    assert len(bad) == 0, f"first mismatch (stream, frame) = {bad[0]}"


@pytest.mark.parametrize("fs,dms,br", [(48000, 100, 128000), (48000, 100, 64000), (16000, 100, 32000), (48000, 100, 240000),
                                       (24000, 50, 64000), (48000, 25, 256000), (44100, 100, 128000)])
def test_bodies_error_protection_equals_oracle(L, fs, dms, br):
    """EP modes 1-4 (lc3_fec.cuh, the concealment-block tracking in lc3_bits.cuh) against the oracle's lc3o_fec.c."""
    N = (48000 if fs == 44100 else fs) // 100 * dms // 100
    nf, S = 10, 5
    pcm = make_streams(S, nf * N, fs, N, 16, silence_frames=2, first_stream=br // 1000)
    for ep in (1, 2, 3, 4):
        epmr = (ep + br // 1000) % 4
        a = hs.sim_encode(L, fs, dms, [(br, nf)], pcm, 16, ep=(ep, epmr))[0]
        b = hs.oracle_encode(fs, dms, [(br, nf)], pcm, 16, ep=(ep, epmr))[0]
        assert a.shape == b.shape
        assert np.array_equal(a, b), (ep, epmr, np.where((a != b).any(axis=2)))
