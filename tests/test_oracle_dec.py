"""The decoder oracle (oracle/lc3o_dec.c, plain-C restatement of fx/dec_lc3.c for good frames) is PINNED: sample for sample
against the unmodified reference decoder compiled from /root/reference (oracle/_ref/liblc3ref.so), and against the committed
digests of that decoder's output (tests/golden/dec_golden.json) where the reference library is absent."""
import hashlib
import json
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import lc3ref  # noqa: E402
from make_golden_dec import CASES, payloads  # noqa: E402
from audio_codec_b200.synth import make_streams, make_bursts  # noqa: E402


@pytest.fixture(scope="module")
def golden():
    return json.load(open(os.path.join(HERE, "golden", "dec_golden.json")))


@pytest.mark.parametrize("case", CASES, ids=lambda c: "%d-%d-%d-%d" % c[:4])
def test_oracle_decoder_equals_reference_digests(golden, case):
    fs, dms, br, bits, S, nf = case
    frames = payloads(case)
    key = "%d-%d-%d-%d" % case[:4]
    assert hashlib.md5(frames.tobytes()).hexdigest() == golden[key]["payload_md5"]
    pcm = []
    for s in range(S):
        o, bad = lc3ref.OracleDec(fs, 1, dms).decode(frames[s], bits)
        assert bad == 0
        pcm.append(o[0])
    assert hashlib.md5(np.ascontiguousarray(np.stack(pcm)).tobytes()).hexdigest() == golden[key]["pcm_md5"]


@pytest.mark.parametrize("fs,dms,rates,ch", [(48000, 100, (32000, 128000, 48000, 320000), 1), (16000, 100, (16000, 64000, 24000), 1),
                                             (32000, 50, (40000, 200000, 64000), 2), (8000, 25, (64000, 100000), 1), (44100, 100, (64000, 32000), 2),
                                             (24000, 25, (96000, 64000), 1), (48000, 50, (64000, 250000), 1)])
def test_oracle_decoder_equals_reference_with_switching_and_stereo(fs, dms, rates, ch):
    """long-ish streams with pitch changes and onsets, the payload size changing between segments, mono and stereo, 16 and 24 bit"""
    if not lc3ref.have_ref():
        pytest.skip("reference library absent")
    N = (48000 if fs == 44100 else fs) // 100 * dms // 100
    seg = 25 * (100 // dms)
    pcm = np.concatenate([make_streams(ch, seg * len(rates) * N // 2, fs, N, 16, first_stream=9, silence_frames=4),
                          make_bursts(ch, seg * len(rates) * N - seg * len(rates) * N // 2, fs, 16, seed=fs // 100 + dms)], axis=1)
    enc = lc3ref.RefEnc(fs, ch, dms, ch * rates[0])
    for bits in (16, 24):
        ref, ora = lc3ref.RefDec(fs, ch, dms), lc3ref.OracleDec(fs, ch, dms)
        for k, br in enumerate(rates):
            assert enc.set_bitrate(ch * br) == 0
            frames = enc.encode(pcm[:, k * seg * N:(k + 1) * seg * N])
            a, bad_a = ref.decode(frames, bits)
            b, bad_b = ora.decode(frames, bits)
            assert bad_a == 0 and bad_b == 0 and np.array_equal(a, b), (bits, k, br)


@pytest.mark.parametrize("item,fs,br", [("thetest8", 8000, 32000), ("thetest16", 16000, 32000), ("thetest24", 24000, 48000),
                                        ("thetest32", 32000, 48000), ("thetest44", 44100, 64000), ("thetest48", 48000, 64000)])
def test_oracle_decoder_equals_reference_on_etsi_items(item, fs, br):
    """the ETSI package's own test items (speech and music) at the package's bitrates, no frame erasures"""
    if not lc3ref.have_ref():
        pytest.skip("reference library absent")
    z = np.load(os.path.join(HERE, "golden", "etsi_full.npz"))
    N = (48000 if fs == 44100 else fs) // 100
    nf = min(len(z[item]) // N, 500)
    frames = lc3ref.RefEnc(fs, 1, 100, br).encode(np.ascontiguousarray(z[item][:nf * N][None, :]))
    for bits in (16, 24):
        a, bad_a = lc3ref.RefDec(fs, 1, 100).decode(frames, bits)
        b, bad_b = lc3ref.OracleDec(fs, 1, 100).decode(frames, bits)
        assert bad_a == 0 and bad_b == 0 and np.array_equal(a, b), bits


def test_oracle_decoder_flags_bad_frames_like_the_reference():
    if not lc3ref.have_ref():
        pytest.skip("reference library absent")
    fs, dms, br, S, nf = 32000, 100, 64000, 5, 40
    frames = payloads((fs, dms, br, 16, S, nf)).copy()
    rng = np.random.RandomState(11)
    for s in range(S):
        for f in range(4, nf, 5):
            frames[s, f, rng.randint(0, frames.shape[2], 6)] ^= rng.randint(1, 256, 6).astype(np.uint8)
    total = 0
    for s in range(S):
        ca, cb = [], []
        a, _ = lc3ref.RefDec(fs, 1, dms).decode(frames[s], 16, ca)
        b, bad = lc3ref.OracleDec(fs, 1, dms).decode(frames[s], 16, cb)
        assert [int(c != 0) for c in ca] == [int(c != 0) for c in cb]
        first = ca.index(2) if 2 in ca else nf
        assert np.array_equal(a[0][:first * 320], b[0][:first * 320])
        total += bad
    assert total > 0
