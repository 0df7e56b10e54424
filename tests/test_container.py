"""Bitstream containers: the library's .lc3plus / G.192 writers and switching-file reader against
the files the reference tool itself writes (fx/codec_exe.c:726-751, :799-843, :968-977; ETSI/Readme.txt:308-333).

CPU part: the reference CLI (oracle/_ref/LC3plus) encodes a WAV; its file is taken apart and re-packed by the library
-- header, frames and, for G.192, the .cfg side file must come back byte for byte.  GPU part: PCM -> lc3plus_enc_batch
-> lc3plus_write_bitstream_file must equal the file of the reference CLI run on the same PCM.
"""
import os
import struct
import subprocess

import numpy as np
import pytest

import audio_codec_b200 as ac
from test_cli_dropin import REF, write_wav

needs_ref = pytest.mark.skipif(not os.path.exists(REF), reason="oracle/_ref/LC3plus not built")


def run_ref(tmp_path, fs, ch, ms, rate, extra=(), n=None, bits=16):
    wav, out = str(tmp_path / "in.wav"), str(tmp_path / ("ref.g192" if "-formatG192" in extra else "ref.lc3"))
    write_wav(wav, fs, ch, bits, n or fs // 4, seed=fs + ch)
    r = subprocess.run([REF, "-E", "-q", "-frame_ms", ms, *extra, wav, out, str(rate)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:]
    return wav, out


def split_lc3plus(blob):
    hdr = struct.unpack("<9H", blob[:18])
    assert hdr[0] == 0xcc1c and hdr[1] == 18
    frames, p = [], 18
    while p < len(blob):
        (n,) = struct.unpack_from("<H", blob, p)
        frames.append(blob[p + 2:p + 2 + n])
        p += 2 + n
    return hdr, frames


def test_header_layout():
    h = ac.container_header(44100, 128000, 2, 25, 0, 0x12345)
    assert h == struct.pack("<9H", 0xcc1c, 18, 441, 1280, 2, 250, 0, 0x2345, 0x1)
    assert len(ac.container_frames(np.zeros((0, 40), np.uint8))) == 0
    # ragged sizes, and the G.192 bit order (LSB of each byte first)
    fr = np.array([[0x81, 0xff, 0x00], [0x01, 0x02, 0x03]], np.uint8)
    b = ac.container_frames(fr, num_bytes=[3, 1])
    assert b == bytes([3, 0, 0x81, 0xff, 0x00, 1, 0, 0x01])
    g = np.frombuffer(ac.container_frames(fr, num_bytes=[1, 0], g192=True), dtype="<u2")
    assert list(g) == [0x6B21, 8, 0x81, 0x7f, 0x7f, 0x7f, 0x7f, 0x7f, 0x7f, 0x81, 0x6B21, 0]


def test_switching_file(tmp_path):
    p = tmp_path / "sw.dat"
    vals = np.array([64000, 32000, -5, 2 ** 40], dtype="<i8")
    p.write_bytes(vals.tobytes() + b"\x01\x02")            # trailing partial word is ignored, as fread does
    assert list(ac.read_switching_file(str(p))) == list(vals)
    with pytest.raises(OSError):
        ac.read_switching_file(str(tmp_path / "missing"))


@needs_ref
@pytest.mark.parametrize("fs,ch,ms,rate", [(48000, 1, "10", 128000), (48000, 2, "2.5", 256000), (16000, 1, "5", 32000)])
def test_repack_reference_file(tmp_path, fs, ch, ms, rate):
    _, out = run_ref(tmp_path, fs, ch, ms, rate)
    blob = open(out, "rb").read()
    hdr, frames = split_lc3plus(blob)
    nb = len(frames[0])
    arr = np.frombuffer(b"".join(frames), np.uint8).reshape(len(frames), nb)
    mine = str(tmp_path / "mine.lc3")
    ac.write_bitstream_file(mine, fs, rate, ch, int(float(ms) * 10), hdr[7] | hdr[8] << 16, arr)
    assert open(mine, "rb").read() == blob
    # G.192: the reference tool writes frames and .cfg separately
    _, g = run_ref(tmp_path, fs, ch, ms, rate, extra=("-formatG192",))
    mine_g = str(tmp_path / "mine.g192")
    ac.write_bitstream_file(mine_g, fs, rate, ch, int(float(ms) * 10), hdr[7] | hdr[8] << 16, arr, g192=True)
    assert open(mine_g, "rb").read() == open(g, "rb").read()
    assert open(str(tmp_path / "mine.g192.cfg"), "rb").read() == open(str(tmp_path / "ref.g192.cfg"), "rb").read()


@pytest.mark.gpu
@needs_ref
@pytest.mark.parametrize("fs,ch,ms,rate", [(48000, 1, "10", 128000), (48000, 2, "2.5", 256000), (24000, 1, "5", 48000)])
def test_encode_and_write_equals_reference_tool(tmp_path, fs, ch, ms, rate):
    import wave
    wav, out = run_ref(tmp_path, fs, ch, ms, rate, n=fs // 3 + 17)      # last frame is zero-padded by the tool
    with wave.open(wav, "rb") as w:
        n = w.getnframes()
        inter = np.frombuffer(w.readframes(n), dtype="<i2").reshape(n, ch)
    dms = int(float(ms) * 10)
    enc = ac.Encoder(fs, ch, dms, rate)
    N, nb = enc.frame_len, enc.num_bytes
    F = -(-n // N)
    pcm = np.zeros((1, ch, F * N), np.int16)
    pcm[0, :, :n] = inter.T
    frames = np.zeros((1, F, nb), np.uint8)
    ac.enc_batch([enc], F, pcm, ch * F * N * 2, 16, frames, F * nb)
    mine = str(tmp_path / "mine.lc3")
    ac.write_bitstream_file(mine, fs, rate, ch, dms, n, frames[0])
    assert open(mine, "rb").read() == open(out, "rb").read()


@pytest.mark.gpu
@needs_ref
def test_bitrate_switching_file_equals_reference_tool(tmp_path):
    """-swf: one int64 per frame, per-channel bitrate, wrapping (fx/codec_exe.c:340-349)."""
    import wave
    fs, ch, dms = 32000, 1, 100
    sw = str(tmp_path / "sw.dat")
    np.array([32000, 32000, 64000, 24000, 96000], dtype="<i8").tofile(sw)
    wav, out = str(tmp_path / "in.wav"), str(tmp_path / "ref.lc3")
    write_wav(wav, fs, ch, 16, fs // 4, seed=5)
    r = subprocess.run([REF, "-E", "-q", "-swf", sw, wav, out, "64000"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:]
    with wave.open(wav, "rb") as w:
        n = w.getnframes()
        x = np.frombuffer(w.readframes(n), dtype="<i2")
    vals = ac.read_switching_file(sw)
    enc = ac.Encoder(fs, ch, dms)
    N = fs // 100
    F = -(-n // N)
    pcm = np.zeros(F * N, np.int16)
    pcm[:n] = x
    frames = np.zeros((F, 400), np.uint8)
    sizes = []
    for f in range(F):
        enc.set_bitrate(int(vals[f % len(vals)]) * ch)
        b = enc.enc(pcm[f * N:(f + 1) * N])
        frames[f, :len(b)] = b
        sizes.append(len(b))
    mine = str(tmp_path / "mine.lc3")
    ac.write_bitstream_file(mine, fs, 64000, ch, dms, n, frames, num_bytes=sizes)
    assert open(mine, "rb").read() == open(out, "rb").read()
