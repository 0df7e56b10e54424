"""Decoder kernel bodies compiled for the host (tests/hostsim, test tooling) against the decoder oracle (oracle/lc3o_dec.c),
the committed digests of the reference decoder's output (tests/golden/dec_golden.json, made by tests/golden/make_golden_dec.py)
and, where it is present, the compiled reference decoder itself (oracle/_ref/liblc3ref.so)."""
import hashlib
import json
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "hostsim"))
sys.path.insert(0, os.path.join(HERE, "golden"))
import run_dec as hd  # noqa: E402
import lc3ref  # noqa: E402
from make_golden_dec import CASES, payloads  # noqa: E402


@pytest.fixture(scope="module")
def L():
    return hd.lib()


@pytest.fixture(scope="module")
def golden():
    return json.load(open(os.path.join(HERE, "golden", "dec_golden.json")))


@pytest.mark.parametrize("case", CASES, ids=lambda c: "%d-%d-%d-%d" % c[:4])
def test_bodies_equal_reference_decoder(L, golden, case):
    fs, dms, br, bits, S, nf = case
    frames = payloads(case)
    pcm, bfi = hd.sim_decode(L, fs, dms, frames, bits, cuts=(3, nf // 2))
    key = "%d-%d-%d-%d" % case[:4]
    assert hashlib.md5(frames.tobytes()).hexdigest() == golden[key]["payload_md5"]   # the fixture's input is what was decoded
    assert hashlib.md5(pcm.tobytes()).hexdigest() == golden[key]["pcm_md5"]
    assert int(bfi.sum()) == 0
    for s in range(S):
        ora, bad = lc3ref.OracleDec(fs, 1, dms).decode(frames[s], bits)
        assert bad == 0 and np.array_equal(ora[0], pcm[s]), (key, s)
    if lc3ref.have_ref():
        for s in range(S):
            ref, bad = lc3ref.RefDec(fs, 1, dms).decode(frames[s], bits)
            assert bad == 0 and np.array_equal(ref[0], pcm[s]), (key, s)


def test_bad_frames_are_flagged_like_the_reference(L):
    Dec = lc3ref.RefDec if lc3ref.have_ref() else lc3ref.OracleDec
    fs, dms, br, bits, S, nf = 32000, 100, 64000, 16, 6, 40
    frames = payloads((fs, dms, br, bits, S, nf)).copy()
    rng = np.random.RandomState(7)
    for s in range(S):
        for f in range(5, nf, 7):
            frames[s, f, rng.randint(0, frames.shape[2], 6)] ^= rng.randint(1, 256, 6).astype(np.uint8)
    pcm, bfi = hd.sim_decode(L, fs, dms, frames, bits)
    n_bad = 0
    for s in range(S):
        codes = []
        ref, bad = Dec(fs, 1, dms).decode(frames[s], bits, codes)
        assert [int(c != 0) for c in codes] == [int(v) for v in bfi[:, s]], s
        n_bad += bad
        first = codes.index(2) if 2 in codes else nf   # identical up to the first concealed frame
        N = fs // 100
        assert np.array_equal(ref[0][:first * N], pcm[s][:first * N])
    assert n_bad > 0
