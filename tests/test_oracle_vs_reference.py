"""Pins the oracle (oracle/*.c) against the UNMODIFIED reference compiled from /root/reference
(oracle/_ref/liblc3ref.so) and against the reference's own known-answer vectors
(ETSI testvec/md5_bin.txt, all twelve lines: EP0 and EP4; hashes are stored without their first hex digit,
testvecCheck.pl:212).  Needs /root/reference (this container); skipped on the GPU box."""
import hashlib
import os
import struct
import wave

import numpy as np
import pytest

import lc3ref
from audio_codec_b200.synth import make_streams, make_torture

ETSI = "/root/reference/LC3plus_ETSI_src_v17171_20200723"
needs_ref = pytest.mark.skipif(not (lc3ref.have_ref() and os.path.isdir(ETSI)), reason="reference tree absent")

VEC = [(8, 32000), (16, 32000), (24, 48000), (32, 48000), (44, 64000), (48, 64000)]


def _read_wav(path):
    with wave.open(path, "rb") as w:
        assert w.getnchannels() == 1 and w.getsampwidth() == 2
        return w.getframerate(), np.frombuffer(w.readframes(w.getnframes()), dtype=np.int16)


@needs_ref
@pytest.mark.parametrize("ep", [0, 4])
@pytest.mark.parametrize("tag,br", VEC)
def test_etsi_md5_vectors(tag, br, ep):
    fs, x = _read_wav(f"{ETSI}/testvec/input/thetest{tag}.wav")
    md5 = {}
    for line in open(f"{ETSI}/testvec/md5_bin.txt"):
        k, v = line.strip().split(":")
        md5[k] = v
    want = md5[f"thetest{tag}_{br}_EP{ep}.lc3plus"]
    enc = lc3ref.OracleEnc(fs, 1, 100)
    assert enc.set_ep_mode(ep) == 0 and enc.set_bitrate(br) == 0   # testvecCheck.pl:107 `-epmode $ep`
    N = enc.frame_len
    nf = (len(x) + N - 1) // N
    xp = np.zeros(nf * N, dtype=np.int16)
    xp[:len(x)] = x
    frames = enc.encode(xp)
    # header fields: fx/codec_exe.c:736-746 (frame_ms*100 = dms*10)
    hdr = struct.pack("<9H", 0xcc1c, 18, fs // 100, br // 100, 1, 100 * 10, ep, len(x) & 0xffff, len(x) >> 16)
    blob = hdr + b"".join(struct.pack("<H", frames.shape[1]) + f.tobytes() for f in frames)
    got = hashlib.md5(blob).hexdigest()
    assert got[1:] == want, (got, want)


EP_CASES = [
    # fs, ch, dms, bitrate: slots of 40 (special CRC), 60, 80+ (partial concealment in modes 3/4), 300 bytes; stereo with
    # combined channel coding (<= 160 bytes per frame) and with one slot per channel
    (16000, 1, 100, 32000), (48000, 1, 100, 48000), (48000, 1, 100, 64000), (48000, 1, 100, 128000), (48000, 1, 100, 240000),
    (48000, 2, 100, 96000), (48000, 2, 100, 128000), (48000, 2, 100, 256000), (24000, 1, 50, 64000), (48000, 1, 25, 256000),
    (44100, 1, 100, 128000), (32000, 1, 100, 96000), (48000, 1, 50, 128000),
]


@needs_ref
@pytest.mark.parametrize("fs,ch,dms,br", EP_CASES)
def test_oracle_error_protection_equals_reference(fs, ch, dms, br):
    """fx/al_fec.c fec_encoder + fx/reorder_bitstream_fx.c + the encoder-side decoder pass (fx/enc_lc3.c:247-312)."""
    ran = 0
    for ep in (1, 2, 3, 4):
        for epmr in (0, 1, 2, 3):
            r, o = lc3ref.RefEnc(fs, ch, dms), lc3ref.OracleEnc(fs, ch, dms)
            assert r.set_ep_mode(ep) == o.set_ep_mode(ep) == 0
            rr, ro = r.set_bitrate(br), o.set_bitrate(br)
            assert rr == ro
            if rr:
                continue
            assert r.set_ep_mode_request(epmr) == o.set_ep_mode_request(epmr) == 0
            N = r.frame_len
            assert r.num_bytes == o.num_bytes
            nf = 24 if epmr in (0, 3) else 6
            pcm = make_streams(ch, nf * N, fs, N, first_stream=ep + br // 1000, silence_frames=3)
            assert np.array_equal(r.encode(pcm), o.encode(pcm)), (ep, epmr)
            ran += 1
    assert ran > 0


@needs_ref
def test_oracle_error_protection_switching_and_errors():
    """ep mode changed between frames (fx/codec_exe.c:351-363) re-derives the payload size; bad arguments."""
    fs, N = 48000, 480
    r, o = lc3ref.RefEnc(fs, 1, 100, 128000), lc3ref.OracleEnc(fs, 1, 100, 128000)
    pcm = make_streams(1, 40 * N, fs, N, first_stream=9)
    for f in range(40):
        ep = (0, 4, 3, 1, 2, 4)[(f // 3) % 6]
        assert r.set_ep_mode(ep) == o.set_ep_mode(ep) == 0
        a, b = r.encode(pcm[:, f * N:(f + 1) * N]), o.encode(pcm[:, f * N:(f + 1) * N])
        assert np.array_equal(a, b), (f, ep)
    assert r.set_ep_mode(5) == o.set_ep_mode(5) == 9
    assert r.set_ep_mode_request(4) == o.set_ep_mode_request(4) == 12
    r2, o2 = lc3ref.RefEnc(16000, 1, 100), lc3ref.OracleEnc(16000, 1, 100)
    assert r2.set_ep_mode(3) == o2.set_ep_mode(3) == 0
    assert r2.set_bitrate(16000) == o2.set_bitrate(16000) == 6          # 20-byte slot < FEC_SLOT_BYTES_MIN


CASES = [
    # fs, ch, dms, bitrate, bits
    (48000, 1, 100, 128000, 16), (48000, 1, 100, 64000, 16), (48000, 1, 100, 320000, 16),
    (16000, 1, 50, 32000, 16), (48000, 2, 25, 256000, 16), (8000, 1, 25, 64000, 16),
    (24000, 1, 50, 48000, 16), (32000, 1, 100, 80000, 16), (44100, 1, 100, 96000, 16),
    (44100, 1, 50, 128000, 16), (32000, 2, 25, 192000, 16), (8000, 1, 100, 16000, 16),
    (48000, 1, 100, 128000, 24), (16000, 1, 50, 32000, 24), (48000, 2, 25, 256000, 24),
    (24000, 1, 100, 16000, 16), (48000, 1, 50, 48000, 16),
]


@needs_ref
@pytest.mark.parametrize("fs,ch,dms,br,bits", CASES)
def test_oracle_equals_reference(fs, ch, dms, br, bits):
    ref = lc3ref.RefEnc(fs, ch, dms, br)
    ora = lc3ref.OracleEnc(fs, ch, dms, br)
    N = ref.frame_len
    assert ora.frame_len == N and ora.num_bytes == ref.num_bytes
    nf = 60
    for s0 in (0, 7, 13):  # stream 7 starts with digital silence
        pcm = make_streams(ch, nf * N, fs, N, bits, first_stream=s0, silence_frames=20)
        a = ref.encode(pcm, bits)
        b = ora.encode(pcm, bits)
        assert np.array_equal(a, b), f"first differing frame {np.nonzero((a != b).any(1))[0][:1]}"


@needs_ref
def test_oracle_bitrate_switching_and_bfi():
    fs, dms = 48000, 100
    ref = lc3ref.RefEnc(fs, 1, dms, 128000)
    ora = lc3ref.OracleEnc(fs, 1, dms, 128000)
    N = ref.frame_len
    pcm = make_streams(1, 40 * N, fs, N, 16, first_stream=3)
    for k, br in enumerate([128000, 32000, 320000, 96000, 64000]):
        assert ref.set_bitrate(br) == ora.set_bitrate(br) == 0
        seg = pcm[:, k * 8 * N:(k + 1) * 8 * N]
        bfi = [0, 0, 1, 0, 0, 0, 1, 0]
        assert np.array_equal(ref.encode(seg, 16, bfi), ora.encode(seg, 16, bfi))


@needs_ref
def test_oracle_error_codes():
    R, O = lc3ref.ref_lib(), lc3ref.oracle_lib()
    ref = lc3ref.RefEnc(16000, 1, 100)
    ora = lc3ref.OracleEnc(16000, 1, 100)
    for br in (1000, 15999, 16000, 320000, 320001, 10**6):
        assert ref.set_bitrate(br) == ora.set_bitrate(br), br
    assert R.lc3_enc_set_frame_dms(ref.h, 50) == O.lc3o_enc_set_frame_dms(ora.h, 50) == 15
    assert O.lc3o_enc_set_frame_dms(ora.h, 30) == 10


@needs_ref
@pytest.mark.parametrize("fs,dms,br,bits", [(48000, 100, 128000, 16), (48000, 100, 320000, 24), (16000, 50, 32000, 16),
                                            (48000, 25, 256000, 16), (32000, 100, 64000, 16), (8000, 100, 16000, 24)])
def test_oracle_equals_reference_on_adversarial_signals(fs, dms, br, bits):
    N = (48000 if fs == 44100 else fs) // 100 * dms // 100
    pcm = make_torture(20 * N, fs, bits)
    for s in range(pcm.shape[0]):
        a = lc3ref.RefEnc(fs, 1, dms, br).encode(pcm[s:s + 1], bits)
        b = lc3ref.OracleEnc(fs, 1, dms, br).encode(pcm[s:s + 1], bits)
        assert np.array_equal(a, b), f"stream {s}: first differing frame {np.nonzero((a != b).any(1))[0][:1]}"
