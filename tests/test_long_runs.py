"""Parity at the durations and sizes that are timed:
  * the six full-length ETSI items through lc3plus_enc_batch + lc3plus_write_bitstream_file, MD5 of the file equal to
    ETSI/testvec/md5_bin.txt (EP0 and EP4) -- the reference's own known answers (tests/golden/etsi_full.npz holds the WAV samples);
  * >= 2000-frame runs of C1 / C3 / C4 against digests of the unmodified reference (tests/golden/make_golden_long.py);
  * full-size C3 (12 500 streams) and C4 (8192 stereo encoders) batches: samples of streams equal the oracle, duplicates equal;
  * C5 as SURVEY.md 8(d) defines it: 540 streams round-robin over {8, 24, 32, 44.1, 48 kHz} x {2.5, 5, 10 ms}, a new legal
    bitrate every 50 frames per stream (seed 99 + s) through lc3_enc_set_bitrate between batches, several chunks per call.
CPU (not gpu): the oracle on one long case, so the digests are pinned on both sides."""
import hashlib
import importlib.util
import json
import os
import tempfile

import numpy as np
import pytest

import lc3ref
from audio_codec_b200.synth import make_streams, make_streams_fast

HERE = os.path.dirname(os.path.abspath(__file__))
spec = importlib.util.spec_from_file_location("make_golden_long", os.path.join(HERE, "golden", "make_golden_long.py"))
ml = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ml)
LONG = json.load(open(os.path.join(HERE, "golden", "long_runs.json")))
# ETSI/testvec/md5_bin.txt:1-12 (stored without the first hex digit, testvecCheck.pl:212)
ETSI_MD5 = {(8, 32000, 0): "40088b4fd8f53a204eccc20c6ad7da0", (8, 32000, 4): "ea2e3f2eb31a8081701ebb73a6266f7",
            (16, 32000, 0): "a433cc05245974eaffc89e8a6c16c4f", (16, 32000, 4): "c0a31e8351d269f492c136b6d8bc350",
            (24, 48000, 0): "faf2758d4f7f9899ab573eeb2f194e4", (24, 48000, 4): "6b7e11b3266e40cbafb144b9c21ed1c",
            (32, 48000, 0): "d546537162ed22042797e320e02260a", (32, 48000, 4): "88ef4b67aaf717984c621ed7aebe9cd",
            (44, 64000, 0): "3e1a5cfb7ba5243cbabbbd53aa3ccb1", (44, 64000, 4): "26d63c8e74ab09e7188fed89e727253",
            (48, 64000, 0): "97bfe0d6bfde9b6fc3f44ffeb432c6f", (48, 64000, 4): "2aaa581fb94d42a9d42993dcb9f1f5d"}
FS_OF = {8: 8000, 16: 16000, 24: 24000, 32: 32000, 44: 44100, 48: 48000}


def _frame_len(fs, dms):
    return (48000 if fs == 44100 else fs) // 100 * dms // 100


def test_oracle_long_run_matches_reference_digest():
    case = ml.LONG_CASES[1]   # C3, 2400 frames
    name, fs, ch, dms, br, bits, nf = case
    got = hashlib.sha256(lc3ref.OracleEnc(fs, ch, dms, br).encode(ml.long_pcm(case, 7), bits).tobytes()).hexdigest()
    assert got == LONG[f"{name}_s7"]


@pytest.mark.gpu
@pytest.mark.parametrize("ep", [0, 4])
def test_cuda_full_etsi_items_md5(ep):
    """All six items in ONE batch call (six different sample rates = six buckets), then the reference tool's file format."""
    import audio_codec_b200 as ac
    z = np.load(os.path.join(HERE, "golden", "etsi_full.npz"))
    items = [(tag, br) for (tag, br, e) in ETSI_MD5 if e == ep]
    encs, xs, nfs = [], [], []
    for tag, br in items:
        fs = FS_OF[tag]
        e = ac.Encoder(fs, 1, 100)
        assert e.set_ep_mode(ep) == 0
        e.set_bitrate(br)
        x = z[f"thetest{tag}"]
        encs.append(e); xs.append(x); nfs.append((len(x) + e.frame_len - 1) // e.frame_len)
    nf = max(nfs)   # 10 s items: 1009 frames each; pad to a common count, the extra frames are cut below
    stride = nf * 480 * 2
    buf = np.zeros((len(items), stride // 2), dtype=np.int16)
    for k, x in enumerate(xs):
        buf[k, :len(x)] = x
    ostride = nf * max(e.num_bytes for e in encs)
    out = np.zeros((len(items), ostride), dtype=np.uint8)
    ac.enc_batch(encs, nf, buf, stride, 16, out, ostride)
    for k, (tag, br) in enumerate(items):
        nb = encs[k].num_bytes
        frames = out[k, :nfs[k] * nb].reshape(nfs[k], nb)
        with tempfile.TemporaryDirectory() as d:
            p = os.path.join(d, "o.lc3plus")
            ac.write_bitstream_file(p, FS_OF[tag], br, 1, 100, len(xs[k]), frames, epmode=ep)
            got = hashlib.md5(open(p, "rb").read()).hexdigest()
        assert got[1:] == ETSI_MD5[(tag, br, ep)], (tag, br, ep)


@pytest.mark.gpu
@pytest.mark.parametrize("case", ml.LONG_CASES, ids=lambda c: c[0])
def test_cuda_long_runs_match_reference_digests(case):
    """2000+ frames per stream in one call (the rate-loop feedback, LTPF hysteresis and the plain-arithmetic-with-fallback
    paths need long signals); host buffers, so the chunked copy pipeline runs too."""
    import audio_codec_b200 as ac
    name, fs, ch, dms, br, bits, nf = case
    pcm = np.stack([ml.long_pcm(case, s0) for s0 in ml.STREAMS])
    encs = [ac.Encoder(fs, ch, dms, br) for _ in ml.STREAMS]
    nb = encs[0].num_bytes
    out = np.zeros((len(encs), nf, nb), dtype=np.uint8)
    ac.set_chunk_frames(32 * 256)     # several chunks of 256 frames
    try:
        ac.enc_batch(encs, nf, pcm, pcm[0].nbytes, bits, out, nf * nb)
    finally:
        ac.set_chunk_frames(0)
    for k, s0 in enumerate(ml.STREAMS):
        if hashlib.sha256(out[k].tobytes()).hexdigest() != LONG[f"{name}_s{s0}"]:
            want = lc3ref.OracleEnc(fs, ch, dms, br).encode(pcm[k], bits)
            bad = np.nonzero((want != out[k]).any(axis=1))[0]
            pytest.fail(f"{name} stream {s0}: first differing frame {bad[:5]} of {nf}")


def _sample_equals_oracle(out, pcm, fs, ch, dms, br, idx):
    for s in idx:
        want = lc3ref.OracleEnc(fs, ch, dms, br).encode(pcm[s].reshape(ch, -1))
        assert np.array_equal(out[s], want), s


@pytest.mark.gpu
def test_full_size_c3_batch():
    """BASELINE configs[2] per GPU: 12 500 streams x 16 kHz / 5 ms / 32 kbps x 200 frames (391 tiles, the last one ragged)."""
    import torch
    import audio_codec_b200 as ac
    fs, dms, br, S, nf = 16000, 50, 32000, 12500, 200
    N = _frame_len(fs, dms)
    pcm = make_streams_fast(S, nf * N, fs, N, n_unique=64)
    pcm[6400:6464] = pcm[0:64]
    encs = ac.EncoderArray([ac.Encoder(fs, 1, dms, br) for _ in range(S)])
    nb = encs.encoders[0].num_bytes
    d_pcm = torch.from_numpy(pcm).cuda()
    d_out = torch.zeros((S, nf, nb), dtype=torch.uint8, device="cuda")
    ac.enc_batch(encs, nf, d_pcm.data_ptr(), nf * N * 2, 16, d_out.data_ptr(), nf * nb)
    torch.cuda.synchronize()
    out = d_out.cpu().numpy()
    assert np.array_equal(out[6400:6464], out[0:64])
    _sample_equals_oracle(out, pcm, fs, 1, dms, br, (0, 7, 31, 32, 4095, 12479, 12480, 12499))
    for e in encs.encoders:
        e.close()


@pytest.mark.gpu
def test_full_size_c4_batch():
    """BASELINE configs[3]: 8192 stereo encoders x 48 kHz / 2.5 ms / 256 kbps x 400 frames (16 384 lanes), host buffers."""
    import audio_codec_b200 as ac
    fs, dms, br, S, nf = 48000, 25, 256000, 8192, 400
    N = _frame_len(fs, dms)
    mono = make_streams_fast(2 * S, nf * N, fs, N, n_unique=64)
    pcm = mono.reshape(S, 2 * nf * N)                 # encoder s: channel 0 = mono stream 2s, channel 1 = 2s + 1 (planar)
    pcm[4096:4128] = pcm[0:32]
    encs = ac.EncoderArray([ac.Encoder(fs, 2, dms, br) for _ in range(S)])
    nb = encs.encoders[0].num_bytes
    out = np.zeros((S, nf, nb), dtype=np.uint8)
    ac.enc_batch(encs, nf, pcm, 2 * nf * N * 2, 16, out, nf * nb)
    assert np.array_equal(out[4096:4128], out[0:32])
    _sample_equals_oracle(out, pcm, fs, 2, dms, br, (0, 3, 15, 16, 31, 4100, 8191))
    for e in encs.encoders:
        e.close()


def _legal_bitrates(fs, dms, ch=1):
    """8 kbps grid inside the legal range of the frame size (ETSI/Readme.txt:277-284; fx/setup_enc_lc3.c:137-149)."""
    k = (441.0 / 480) if fs == 44100 else 1.0
    lo = int(20 * 8.0 * (10000.0 / dms) * k) * ch
    hi = int(400 * 8.0 * (10000.0 / dms) * k) * ch
    return np.arange((lo + 7999) // 8000 * 8000, hi + 1, 8000)


@pytest.mark.gpu
def test_c5_mixed_batch_with_bitrate_switching():
    import audio_codec_b200 as ac
    FS = (8000, 24000, 32000, 44100, 48000)
    DMS = (25, 50, 100)
    S, seg, nseg = 540, 50, 4
    cfgs = [(FS[s % 5], DMS[(s // 5) % 3]) for s in range(S)]
    rngs = [np.random.RandomState(99 + s) for s in range(S)]
    sched = [[int(rngs[s].choice(_legal_bitrates(*cfgs[s]))) for _ in range(nseg)] for s in range(S)]
    for s in range(0, S, 45):      # some streams cross the attack detector's on/off threshold back and forth (48 kHz, 10 ms)
        if cfgs[s] == (48000, 100):
            sched[s] = [128000, 32000, 128000, 64000]
    encs = ac.EncoderArray([ac.Encoder(fs, 1, dms) for fs, dms in cfgs])
    oras = [lc3ref.OracleEnc(fs, 1, dms) for fs, dms in cfgs]
    pcms = [make_streams(1, seg * nseg * _frame_len(fs, dms), fs, _frame_len(fs, dms), 16, first_stream=s, silence_frames=60 if s % 8 == 7 else 0)[0]
            for s, (fs, dms) in enumerate(cfgs)]
    stride = seg * 480 * 2
    ac.set_chunk_frames(64 * 16)   # 2-tile buckets: 16 frames per chunk, four chunks per call
    try:
        for k in range(nseg):
            buf = np.zeros((S, stride // 2), dtype=np.int16)
            for s, (fs, dms) in enumerate(cfgs):
                N = _frame_len(fs, dms)
                buf[s, :seg * N] = pcms[s][k * seg * N:(k + 1) * seg * N]
                if k == 2 and s % 7 == 0:   # a switch that is overwritten before any frame is coded (state clears must still match)
                    lo = int(_legal_bitrates(fs, dms)[0])
                    assert encs.encoders[s].set_bitrate(lo) == 0 and oras[s].set_bitrate(lo) == 0
                assert encs.encoders[s].set_bitrate(sched[s][k]) == 0 and oras[s].set_bitrate(sched[s][k]) == 0
            ostride = seg * max(e.num_bytes for e in encs.encoders)
            out = np.zeros((S, ostride), dtype=np.uint8)
            ac.enc_batch(encs, seg, buf, stride, 16, out, ostride)
            for s in range(S):
                N = _frame_len(*cfgs[s])
                want = oras[s].encode(buf[s, :seg * N][None, :])
                assert np.array_equal(out[s, :want.size].reshape(want.shape), want), (s, cfgs[s], k, sched[s])
    finally:
        ac.set_chunk_frames(0)
    for e in encs.encoders:
        e.close()


@pytest.mark.gpu
def test_bit_depth_switch_is_refused():
    """One encoder, one bit depth (include/lc3plus_b200.h): the switch returns LC3_ERROR, a fresh encoder works."""
    import audio_codec_b200 as ac
    e = ac.Encoder(48000, 1, 100, 128000)
    x16 = make_streams(1, 480, 48000, 480, 16)
    x24 = make_streams(1, 480, 48000, 480, 24)
    e.enc(x16, 16)
    with pytest.raises(ac.LC3Error) as ex:
        e.enc(x24, 24)
    assert ex.value.code == 1 and "bit depth" in str(ex.value)
    e.enc(x16, 16)


@pytest.mark.gpu
@pytest.mark.parametrize("fs,ch,dms,br,bits,ep,device", [(48000, 2, 25, 256000, 16, 0, False), (24000, 2, 50, 96000, 24, 0, False),
                                                         (48000, 1, 100, 128000, 16, 3, False), (32000, 1, 100, 64000, 16, 0, True),
                                                         (48000, 2, 100, 192000, 24, 4, True), (16000, 1, 25, 64000, 16, 0, True)])
def test_chunked_pipeline_variants(fs, ch, dms, br, bits, ep, device):
    """The multi-chunk pipeline beyond mono/16-bit/N=480: stereo host copies, the 24-bit chain, the short-frame front end,
    lost-frame markers across chunks (frame0 indexing) and error protection per chunk; host and device buffers."""
    import torch
    import audio_codec_b200 as ac
    N, nf, S = _frame_len(fs, dms), 23, 5
    pcm = np.stack([make_streams(ch, nf * N, fs, N, bits, first_stream=3 + s * ch, silence_frames=5) for s in range(S)])
    bfi = (np.arange(S * nf) % 7 == 3)
    encs, oras = [], []
    for s in range(S):
        e, o = ac.Encoder(fs, ch, dms), lc3ref.OracleEnc(fs, ch, dms)
        assert e.set_ep_mode(ep) == o.set_ep_mode(ep) == 0
        assert e.set_bitrate(br) == o.set_bitrate(br) == 0
        encs.append(e); oras.append(o)
    nb = encs[0].num_bytes
    num_bytes = np.where(bfi, -1, 0).astype(np.int32)
    ac.set_chunk_frames(32 * 3)    # 3 frames per chunk: 8 chunks, ramped first chunks with host buffers
    try:
        if device:
            d_pcm = torch.from_numpy(pcm).cuda()
            d_out = torch.zeros((S, nf, nb), dtype=torch.uint8, device="cuda")
            ac.enc_batch(encs, nf, d_pcm.data_ptr(), pcm[0].nbytes, bits, d_out.data_ptr(), nf * nb, num_bytes)
            torch.cuda.synchronize()
            out = d_out.cpu().numpy()
        else:
            out = np.zeros((S, nf, nb), dtype=np.uint8)
            ac.enc_batch(encs, nf, pcm, pcm[0].nbytes, bits, out, nf * nb, num_bytes)
    finally:
        ac.set_chunk_frames(0)
    assert (num_bytes == nb).all()
    for s in range(S):
        want = oras[s].encode(pcm[s], bits, bfi[s * nf:(s + 1) * nf])
        assert np.array_equal(out[s], want), s


@pytest.mark.gpu
@pytest.mark.parametrize("fs,ch,dms,br,ep", [(48000, 1, 100, 128000, 0), (16000, 1, 50, 32000, 0), (48000, 2, 25, 256000, 0),
                                             (48000, 1, 100, 128000, 4), (32000, 2, 100, 128000, 3)])
def test_round_trip_through_the_reference_decoder(fs, ch, dms, br, ep):
    """Size-independent property: what the CUDA encoder writes, the unmodified reference DECODER (oracle/_ref) reads without a
    frame error -- error-protected slots included, whose CRCs and Reed-Solomon words it checks -- and gives back the input
    (delayed by the codec delay) at a signal-to-noise ratio a working codec has at these rates."""
    if not lc3ref.have_ref():
        pytest.skip("oracle/_ref/liblc3ref.so absent")
    import audio_codec_b200 as ac
    N, nf = _frame_len(fs, dms), 1000 // (dms // 25) // 2 + 40
    pcm = make_streams(ch, nf * N, fs, N, 16, first_stream=3)
    e = ac.Encoder(fs, ch, dms)
    assert e.set_ep_mode(ep) == 0
    e.set_bitrate(br)
    nb = e.num_bytes
    out = np.zeros((1, nf, nb), dtype=np.uint8)
    ac.enc_batch([e], nf, pcm[None], pcm.nbytes, 16, out, nf * nb)
    dec = lc3ref.RefDec(fs, ch, dms, ep=bool(ep))
    y, bad = dec.decode(out[0])
    assert bad == 0
    d = dec.delay
    for c in range(ch):
        x = pcm[c, 20 * N:(nf - 2) * N - d].astype(np.float64)
        z = y[c, 20 * N + d:(nf - 2) * N].astype(np.float64)
        snr = 10 * np.log10(np.sum(x * x) / max(np.sum((x - z) ** 2), 1.0))
        assert snr > 12.0, (c, snr)
