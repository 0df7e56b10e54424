"""Makes tests/golden/dec_golden.json: MD5 digests of the reference DECODER's output (oracle/_ref/liblc3ref.so, compiled from
the unmodified sources under /root/reference) for payloads produced by the reference ENCODER from the synthetic streams of
audio_codec_b200/synth.py.  Run in the container that has /root/reference:  python tests/golden/make_golden_dec.py
The tests regenerate the payloads with our own oracle (bit-exact with the reference encoder, checked by payload_md5), so they
run where the reference library is absent."""
import hashlib
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

# (fs, frame_dms, bitrate, output bits, streams, frames)
CASES = [(48000, 100, 128000, 16, 5, 40), (48000, 100, 64000, 24, 4, 40), (16000, 100, 32000, 16, 5, 40), (8000, 100, 24000, 16, 4, 40),
         (24000, 50, 48000, 16, 4, 60), (32000, 100, 64000, 16, 4, 40), (44100, 100, 96000, 16, 4, 40), (48000, 50, 80000, 16, 4, 60),
         (48000, 25, 128000, 16, 4, 80), (8000, 25, 64000, 16, 4, 80), (16000, 50, 32000, 24, 4, 60), (32000, 25, 96000, 16, 4, 80),
         (48000, 100, 320000, 16, 3, 30), (24000, 100, 16000, 16, 4, 40), (16000, 25, 64000, 16, 4, 80), (8000, 50, 32000, 16, 4, 60)]


def payloads(case):
    """uint8 [S, n_frames, nbytes]: the case's streams encoded by the oracle (== the reference encoder, byte for byte)"""
    import lc3ref
    from audio_codec_b200.synth import make_streams, make_bursts
    fs, dms, br, bits, S, nf = case
    N = (48000 if fs == 44100 else fs) // 100 * dms // 100
    pcm = make_streams(S, nf * N, fs, N, 16, first_stream=3, silence_frames=6)
    pcm[S - 1] = make_bursts(1, nf * N, fs, 16, seed=fs // 1000 + dms)[0]   # pitch changes, onsets: LTPF transitions
    out = []
    for s in range(S):
        e = lc3ref.OracleEnc(fs, 1, dms, br)
        out.append(e.encode(pcm[s:s + 1]))
    return np.ascontiguousarray(np.stack(out))


def main():
    import lc3ref
    g = {}
    for case in CASES:
        fs, dms, br, bits, S, nf = case
        frames = payloads(case)
        pcm = []
        for s in range(S):
            o, bad = lc3ref.RefDec(fs, 1, dms).decode(frames[s], bits)
            assert bad == 0
            pcm.append(o[0])
        pcm = np.ascontiguousarray(np.stack(pcm))
        g["%d-%d-%d-%d" % case[:4]] = {"payload_md5": hashlib.md5(frames.tobytes()).hexdigest(),
                                       "pcm_md5": hashlib.md5(pcm.tobytes()).hexdigest(), "streams": S, "frames": nf}
        print(case, g["%d-%d-%d-%d" % case[:4]])
    json.dump(g, open(os.path.join(HERE, "dec_golden.json"), "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
