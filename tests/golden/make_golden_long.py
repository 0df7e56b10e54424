"""Generates tests/golden/long_runs.json: SHA-256 of the UNMODIFIED reference encoder's bitstreams
(oracle/_ref/liblc3ref.so) for >= 2000-frame synthetic runs of the BASELINE configurations C1, C3, C4
(SURVEY.md section 8d).  Run in the container that has /root/reference:  python tests/golden/make_golden_long.py"""
import hashlib
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import lc3ref  # noqa: E402
from audio_codec_b200.synth import make_streams  # noqa: E402

# name, fs, channels, dms, bitrate, bits, n_frames
LONG_CASES = [("c1", 48000, 1, 100, 128000, 16, 2000), ("c3", 16000, 1, 50, 32000, 16, 2400), ("c4", 48000, 2, 25, 256000, 16, 2400),
              ("c1_24bit", 48000, 1, 100, 128000, 24, 2000)]
STREAMS = (0, 7, 21)   # 7: fifty frames of digital silence first


def long_pcm(case, s0):
    name, fs, ch, dms, br, bits, nf = case
    N = fs // 100 * dms // 100
    return make_streams(ch, nf * N, fs, N, bits, first_stream=s0 * ch, silence_frames=50)


def main():
    out = {}
    for case in LONG_CASES:
        name, fs, ch, dms, br, bits, nf = case
        for s0 in STREAMS:
            enc = lc3ref.RefEnc(fs, ch, dms, br)
            out[f"{name}_s{s0}"] = hashlib.sha256(enc.encode(long_pcm(case, s0), bits).tobytes()).hexdigest()
    json.dump(out, open(os.path.join(HERE, "long_runs.json"), "w"), indent=1, sort_keys=True)
    print("wrote", len(out), "digests")


if __name__ == "__main__":
    main()
