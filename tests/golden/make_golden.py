"""Generates tests/golden/bitstreams.json: SHA-256 of the bitstreams the UNMODIFIED reference encoder
(oracle/_ref/liblc3ref.so, compiled from /root/reference by oracle/Makefile) produces for seeded synthetic
inputs (audio_codec_b200.synth.make_streams).  Run in the container that has /root/reference:
    python tests/golden/make_golden.py
The GPU box has no reference tree; there the oracle and the CUDA path are checked against these digests."""
import hashlib
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import lc3ref  # noqa: E402
from audio_codec_b200.synth import make_streams  # noqa: E402

# fs, channels, dms, bitrate, bits, n_frames
CASES = [
    (48000, 1, 100, 128000, 16, 100), (16000, 1, 50, 32000, 16, 100), (48000, 2, 25, 256000, 16, 100),
    (8000, 1, 100, 16000, 16, 50), (8000, 1, 25, 64000, 16, 50), (24000, 1, 50, 48000, 16, 50),
    (32000, 1, 100, 80000, 16, 50), (44100, 1, 100, 96000, 16, 50), (44100, 1, 50, 128000, 16, 50),
    (32000, 2, 25, 192000, 16, 50), (48000, 1, 100, 320000, 16, 50), (48000, 1, 100, 32000, 16, 50),
    (48000, 1, 100, 128000, 24, 50), (16000, 1, 50, 32000, 24, 50), (48000, 2, 25, 256000, 24, 50),
    (24000, 1, 100, 16000, 16, 50), (48000, 1, 50, 48000, 16, 50), (16000, 1, 100, 24000, 16, 50),
    (32000, 1, 50, 64000, 24, 50), (24000, 1, 25, 64000, 16, 50), (16000, 2, 25, 128000, 16, 50),
]
STREAMS = (0, 7, 13)  # stream 7 starts with digital silence
# error protection (fx/al_fec.c): fs, channels, dms, bitrate, ep mode, ep mode request, n_frames.  Slots of 40 / 60 / 80 /
# 160 / 300 bytes (partial concealment from 80 bytes up in modes 3 and 4), stereo jointly coded and per channel
EP_CASES = [
    (16000, 1, 100, 32000, 1, 0, 30), (16000, 1, 100, 32000, 4, 3, 30), (48000, 1, 100, 48000, 2, 1, 30),
    (48000, 1, 100, 64000, 3, 0, 30), (48000, 1, 100, 64000, 4, 2, 30), (48000, 1, 100, 128000, 3, 3, 30),
    (48000, 1, 100, 128000, 4, 0, 30), (48000, 1, 100, 240000, 4, 1, 30), (48000, 2, 100, 96000, 2, 0, 30),
    (48000, 2, 100, 128000, 4, 2, 30), (48000, 2, 100, 256000, 3, 0, 30), (24000, 1, 50, 64000, 4, 0, 30),
    (48000, 1, 25, 256000, 3, 1, 30), (44100, 1, 100, 128000, 4, 0, 30),
]


def ep_key(c):
    return "ep_fs%d_ch%d_dms%d_br%d_ep%d_epmr%d_f%d" % c


def ep_pcm(c, s0):
    fs, ch, dms, br, ep, epmr, nf = c
    N = (48000 if fs == 44100 else fs) // 100 * dms // 100
    return make_streams(ch, nf * N, fs, N, 16, first_stream=s0, silence_frames=5)


def key(c):
    return "fs%d_ch%d_dms%d_br%d_b%d_f%d" % c


def case_pcm(c, s0):
    fs, ch, dms, br, bits, nf = c
    N = (48000 if fs == 44100 else fs) // 100 * dms // 100
    return make_streams(ch, nf * N, fs, N, bits, first_stream=s0, silence_frames=20)


def main():
    out = {}
    for c in CASES:
        fs, ch, dms, br, bits, nf = c
        h = hashlib.sha256()
        for s0 in STREAMS:
            enc = lc3ref.RefEnc(fs, ch, dms, br)
            h.update(enc.encode(case_pcm(c, s0), bits).tobytes())
        out[key(c)] = h.hexdigest()
    for c in EP_CASES:
        fs, ch, dms, br, ep, epmr, nf = c
        h = hashlib.sha256()
        for s0 in STREAMS:
            enc = lc3ref.RefEnc(fs, ch, dms)
            assert enc.set_ep_mode(ep) == 0 and enc.set_bitrate(br) == 0 and enc.set_ep_mode_request(epmr) == 0
            h.update(enc.encode(ep_pcm(c, s0), 16).tobytes())
        out[ep_key(c)] = h.hexdigest()
    # first second of each ETSI test vector (ETSI/testvec/input/thetest*.wav) at the operating point of
    # testvecCheck.pl:41-51, so the GPU box can run real-signal vectors without the reference tree
    import wave
    import numpy as np
    heads = {}
    for tag, br in ((8, 32000), (16, 32000), (24, 48000), (32, 48000), (44, 64000), (48, 64000)):
        with wave.open(f"/root/reference/LC3plus_ETSI_src_v17171_20200723/testvec/input/thetest{tag}.wav", "rb") as w:
            fs = w.getframerate()
            x = np.frombuffer(w.readframes(w.getnframes()), dtype=np.int16)
        N = (48000 if fs == 44100 else fs) // 100
        x = x[200 * N:300 * N].copy()  # one second from the middle of the item
        heads[f"thetest{tag}"] = x
        enc = lc3ref.RefEnc(fs, 1, 100, br)
        out[f"etsi_thetest{tag}_fs{fs}_br{br}"] = hashlib.sha256(enc.encode(x[None, :]).tobytes()).hexdigest()
        enc = lc3ref.RefEnc(fs, 1, 100)  # the EP4 operating point of the same vectors (testvecCheck.pl:107, -epmode 4)
        assert enc.set_ep_mode(4) == 0 and enc.set_bitrate(br) == 0
        out[f"etsi_thetest{tag}_fs{fs}_br{br}_ep4"] = hashlib.sha256(enc.encode(x[None, :]).tobytes()).hexdigest()
    np.savez_compressed(os.path.join(HERE, "etsi_heads.npz"), **heads)
    with open(os.path.join(HERE, "bitstreams.json"), "w") as f:
        json.dump(out, f, indent=1, sort_keys=True)
    print("wrote", len(out), "digests")


if __name__ == "__main__":
    main()
