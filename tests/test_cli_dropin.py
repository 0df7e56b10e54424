"""Drop-in proof through the reference's own command-line tool ("who calls it":
fx/codec_exe.c:192-209 setup, :347 bitrate switch, :369 bandwidth, :418/:424 encode).

oracle/_ref/LC3plus         = the unmodified reference, all of it.
oracle/_ref/LC3plus_dropin  = the same codec_exe.c, source untouched, whose lc3_enc_* AND lc3_dec_* symbols are resolved by
                              audio_codec_b200/liblc3plus_b200.so (first in link order; the rest comes from the reference).
Both encode the same WAV; the .lc3plus files (header + per-frame length + payload, ETSI/Readme.txt:308-333) must be
byte-identical.  Both decode the same .lc3plus file (no error protection: the library has no channel decoder and no
concealment); the WAV files must be byte-identical.  Needs the GPU: the drop-in has no CPU path.
"""
import os
import subprocess
import wave

import numpy as np
import pytest

from audio_codec_b200 import synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref", "LC3plus")
DROPIN = os.path.join(ROOT, "oracle", "_ref", "LC3plus_dropin")

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not (os.path.exists(REF) and os.path.exists(DROPIN)), reason="oracle/_ref CLIs not built")]


def write_wav(path, fs, channels, bits, n_samples, seed):
    pcm = synth.make_streams(channels, n_samples, fs, fs // 100, bits=16, first_stream=seed, silence_frames=3)   # [channels][n] int16
    inter = np.ascontiguousarray(pcm.T)
    with wave.open(path, "wb") as w:
        w.setnchannels(channels)
        w.setframerate(fs)
        if bits == 16:
            w.setsampwidth(2)
            w.writeframes(inter.astype("<i2").tobytes())
        else:
            w.setsampwidth(3)
            v = (inter.astype(np.int32) * 173).astype("<i4")               # spread into 24 bits, not a plain shift
            b = v.view(np.uint8).reshape(-1, 4)[:, :3]
            w.writeframes(b.tobytes())


CASES = [
    # fs, channels, wav bits, frame_ms, bitrate, extra options
    (48000, 1, 16, "10", 128000, []),
    (48000, 2, 16, "2.5", 256000, []),
    (16000, 1, 16, "5", 32000, []),
    (32000, 1, 24, "10", 64000, []),
    (48000, 1, 16, "10", 96000, ["-bandwidth", "8000"]),
    (44100, 2, 16, "10", 160000, []),
    (48000, 1, 16, "10", 128000, ["-epmode", "4"]),     # error protection with partial concealment
    (48000, 2, 16, "10", 96000, ["-epmode", "3"]),      # both channels in one protected slot
    (32000, 1, 16, "10", 64000, ["-epmode", "2"]),
]


@pytest.mark.parametrize("fs,ch,bits,ms,rate,extra", CASES)
def test_reference_cli_on_our_library(tmp_path, fs, ch, bits, ms, rate, extra):
    wav = str(tmp_path / "in.wav")
    write_wav(wav, fs, ch, bits, fs // 2, seed=fs + ch + bits)
    outs = []
    for exe, name in ((REF, "ref.lc3"), (DROPIN, "b200.lc3")):
        out = str(tmp_path / name)
        r = subprocess.run([exe, "-E", "-q", "-frame_ms", ms] + extra + [wav, out, str(rate)],
                           capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, (exe, r.stdout[-2000:], r.stderr[-2000:])
        outs.append(open(out, "rb").read())
    assert len(outs[0]) > 100
    assert outs[0] == outs[1]


@pytest.mark.parametrize("fs,ch,bits,ms,rate,extra", [c for c in CASES if "-epmode" not in c[5]])
def test_reference_cli_decodes_on_our_library(tmp_path, fs, ch, bits, ms, rate, extra):
    """fx/codec_exe.c:206-215 decoder setup, :468/:474 lc3_dec16 / lc3_dec24 per frame, delay compensation and WAV writer of the tool"""
    wav = str(tmp_path / "in.wav")
    write_wav(wav, fs, ch, bits, fs // 2, seed=fs + ch + bits + 1)
    lc3 = str(tmp_path / "ref.lc3")
    r = subprocess.run([REF, "-E", "-q", "-frame_ms", ms] + extra + [wav, lc3, str(rate)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-2000:])
    outs = []
    for exe, name in ((REF, "ref.wav"), (DROPIN, "b200.wav")):
        out = str(tmp_path / name)
        r = subprocess.run([exe, "-D", "-q", "-bps", str(bits), lc3, out], capture_output=True, text=True, timeout=600)
        assert r.returncode == 0, (exe, r.stdout[-2000:], r.stderr[-2000:])
        outs.append(open(out, "rb").read())
    assert len(outs[0]) > 1000
    assert outs[0] == outs[1]
