"""GPU parity of the DECODER through the C ABI: lc3plus_dec_batch / lc3_dec16 / lc3_dec24 against the
compiled reference decoder (oracle/_ref/liblc3ref.so, test infrastructure; where it is absent the decoder oracle, oracle/lc3o_dec.c,
which is pinned on it) and the committed digests of the reference decoder's output.
Error-free streams: the PCM must be identical sample for sample; corrupt payloads: the same frames flagged."""
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
from audio_codec_b200.synth import make_streams, make_streams_fast, make_bursts  # noqa: E402

pytestmark = pytest.mark.gpu


def CheckDec(fs, channels, dms):
    """the checker: the compiled reference decoder where it travels, else the oracle restatement"""
    return lc3ref.RefDec(fs, channels, dms) if lc3ref.have_ref() else lc3ref.OracleDec(fs, channels, dms)


def gpu_encode(fs, dms, br, pcm, channels=1):
    """payloads [S, n_frames, nbytes] from the GPU encoder (bit-exact with the reference encoder: tests/test_gpu_parity.py)"""
    import audio_codec_b200 as ac
    S = pcm.shape[0] // channels
    encs = [ac.Encoder(fs, channels, dms, br) for _ in range(S)]
    N, nb = encs[0].frame_len, encs[0].num_bytes
    nf = pcm.shape[1] // N
    out = np.zeros((S, nf, nb), dtype=np.uint8)
    ac.enc_batch(encs, nf, pcm, channels * nf * N * 2, 16, out, nf * nb)
    for e in encs:
        e.close()
    return out


def gpu_decode(fs, dms, frames, bits=16, channels=1, cuts=(), decs=None, bfi=None):
    import audio_codec_b200 as ac
    frames = np.ascontiguousarray(frames)
    S, nf, nb = frames.shape
    own = decs is None
    if own:
        decs = [ac.Decoder(fs, channels, dms) for _ in range(S)]
    N = decs[0].frame_len
    dt = np.int16 if bits == 16 else np.int32
    pcm = np.zeros((S, channels, nf * N), dtype=dt)
    edges = [0] + list(cuts) + [nf]
    rc = 0
    for a, b in zip(edges[:-1], edges[1:]):
        seg = np.ascontiguousarray(frames[:, a:b])
        o = np.zeros((S, channels, (b - a) * N), dtype=dt)
        fl = np.zeros((S, b - a), dtype=np.uint8) if bfi is not None else None
        rc |= ac.dec_batch(decs, b - a, seg, (b - a) * nb, np.full(S, nb, np.int32), o, channels * (b - a) * N * o.itemsize, bits, fl)
        pcm[:, :, a * N:b * N] = o
        if bfi is not None:
            bfi[:, a:b] = fl
    if own:
        for d in decs:
            d.close()
    return pcm, rc


@pytest.fixture(scope="module")
def golden():
    return json.load(open(os.path.join(HERE, "golden", "dec_golden.json")))


@pytest.mark.parametrize("case", CASES, ids=lambda c: "%d-%d-%d-%d" % c[:4])
def test_dec_batch_equals_reference_digests(golden, case):
    fs, dms, br, bits, S, nf = case
    frames = payloads(case)
    key = "%d-%d-%d-%d" % case[:4]
    assert hashlib.md5(frames.tobytes()).hexdigest() == golden[key]["payload_md5"]
    pcm, rc = gpu_decode(fs, dms, frames, bits, cuts=(3, nf // 2))
    assert rc == 0
    assert hashlib.md5(np.ascontiguousarray(pcm[:, 0]).tobytes()).hexdigest() == golden[key]["pcm_md5"]


@pytest.mark.parametrize("fs,dms,br,bits", [(48000, 100, 128000, 16), (48000, 100, 48000, 24), (16000, 50, 32000, 16), (32000, 25, 96000, 16),
                                            (8000, 100, 16000, 16), (44100, 100, 128000, 16), (24000, 25, 64000, 24), (48000, 50, 160000, 16)])
def test_dec_batch_many_streams_equal_reference(fs, dms, br, bits):
    N = (48000 if fs == 44100 else fs) // 100 * dms // 100
    S, nf = 70, 50 * (100 // dms)
    pcm = make_streams(S, nf * N, fs, N, 16, silence_frames=9)
    pcm[S // 2:] = make_bursts(S - S // 2, nf * N, fs, 16, seed=5)
    frames = gpu_encode(fs, dms, br, pcm)
    out, rc = gpu_decode(fs, dms, frames, bits, cuts=(1, 17))
    assert rc == 0
    for s in range(S):
        ref, bad = CheckDec(fs, 1, dms).decode(frames[s], bits)
        assert bad == 0 and np.array_equal(ref[0], out[s, 0]), s


def test_dec_stereo_and_mixed_byte_counts():
    import audio_codec_b200 as ac
    fs, dms, nf, N = 48000, 100, 30, 480
    rates = [64000, 128000, 255200, 96000]       # 255200: odd byte count, uneven channel split
    S = len(rates)
    pcm = make_streams(2 * S, nf * N, fs, N, 16, silence_frames=4)
    frames, nbs = [], []
    for s, br in enumerate(rates):
        f = gpu_encode(fs, dms, br, pcm[2 * s:2 * s + 2], channels=2)[0]
        frames.append(f); nbs.append(f.shape[1])
    stride = max(nbs) * nf
    buf = np.zeros((S, stride), dtype=np.uint8)
    for s in range(S):
        buf[s, :nf * nbs[s]] = frames[s].reshape(-1)
    decs = [ac.Decoder(fs, 2, dms) for _ in range(S)]
    out = np.zeros((S, 2, nf * N), dtype=np.int16)
    rc = ac.dec_batch(decs, nf, buf, stride, np.array(nbs, np.int32), out, 2 * nf * N * 2, 16)
    assert rc == 0
    for s in range(S):
        ref, bad = CheckDec(fs, 2, dms).decode(frames[s], 16)
        assert bad == 0 and np.array_equal(ref, out[s]), s


def test_dec_mixed_configurations_in_one_call():
    """BASELINE config c5's geometry for the decoder: sample rates, frame durations and byte counts mixed in one call, streams of
    the groups interleaved; also spread over two devices when the box has them (host buffers)"""
    import torch
    import audio_codec_b200 as ac
    cfgs = [(48000, 100, 128000), (16000, 50, 32000), (32000, 25, 96000), (8000, 100, 24000), (24000, 50, 64000)]
    per, ms = 7, 300                       # 300 ms of every stream: 30 / 60 / 120 frames
    frames, meta = [], []
    for k, (fs, dms, br) in enumerate(cfgs):
        N = fs // 100 * dms // 100
        nf = ms * 10 // dms
        pcm = make_streams(per, nf * N, fs, N, 16, first_stream=3 * k, silence_frames=5)
        f = gpu_encode(fs, dms, br, pcm)
        for s in range(per):
            frames.append(f[s]); meta.append((fs, dms, nf, N))
    order = [k * per + s for s in range(per) for k in range(len(cfgs))]      # interleave the groups
    frames = [frames[i] for i in order]; meta = [meta[i] for i in order]
    S = len(frames)
    nf = 30   # a call decodes the same NUMBER of frames of every stream, whatever their duration
    in_stride = max(f.shape[1] for f in frames) * nf
    out_stride = 480 * nf * 2
    buf = np.zeros((S, in_stride), dtype=np.uint8)
    nbs = np.zeros(S, np.int32)
    for i, f in enumerate(frames):
        nbs[i] = f.shape[1]
        buf[i, :nf * f.shape[1]] = f[:nf].reshape(-1)
    ndev = torch.cuda.device_count()
    decs = []
    for i, (fs, dms, _, _) in enumerate(meta):
        ac.set_device(i % 2 if ndev > 1 else -1)
        decs.append(ac.Decoder(fs, 1, dms))
    ac.set_device(-1)
    out = np.zeros((S, out_stride // 2), dtype=np.int16)
    bfi = np.ones((S, nf), dtype=np.uint8)
    rc = ac.dec_batch(decs, nf, buf, in_stride, nbs, out, out_stride, 16, bfi)
    assert rc == 0 and not bfi.any()
    for i, (fs, dms, _, N) in enumerate(meta):
        ref, bad = CheckDec(fs, 1, dms).decode(frames[i][:nf], 16)
        assert bad == 0 and np.array_equal(ref[0], out[i, :nf * N]), (i, fs, dms)
        assert not out[i, nf * N:].any()


@pytest.mark.parametrize("fs,dms,rates", [(16000, 100, (24000, 64000, 16000, 32000)), (48000, 100, (48000, 160000, 32000, 64000)),
                                          (32000, 50, (40000, 128000, 32000)), (24000, 25, (64000, 200000, 96000))])
def test_dec_byte_count_switching_between_calls(fs, dms, rates):
    """the stream's frame size changes from call to call (the encoder switched bitrate, fx/codec_exe.c:340-349): the decoder follows
    update_dec_bitrate (fx/dec_lc3.c:477-482) -- LTPF scale index, gain offset, TNS weighting -- with the LTPF memories carried
    through the change (fade-out with the old index when the new rate switches the filter off)"""
    import audio_codec_b200 as ac
    N = fs // 100 * dms // 100
    S, seg = 12, 20 * (100 // dms)
    total = seg * len(rates)
    pcm = make_streams(S, total * N, fs, N, 16, first_stream=5, silence_frames=3)
    pcm[S // 2:] = make_bursts(S - S // 2, total * N, fs, 16, seed=11)
    encs = [ac.Encoder(fs, 1, dms, rates[0]) for _ in range(S)]
    decs = [ac.Decoder(fs, 1, dms) for _ in range(S)]
    refs = [CheckDec(fs, 1, dms) for _ in range(S)]
    for k, br in enumerate(rates):
        for e in encs:
            e.set_bitrate(br)
        nb = encs[0].num_bytes
        fr = np.zeros((S, seg, nb), dtype=np.uint8)
        ac.enc_batch(encs, seg, np.ascontiguousarray(pcm[:, k * seg * N:(k + 1) * seg * N]), seg * N * 2, 16, fr, seg * nb)
        out = np.zeros((S, seg * N), dtype=np.int16)
        assert ac.dec_batch(decs, seg, fr, seg * nb, np.full(S, nb, np.int32), out, seg * N * 2, 16) == 0
        for s in range(S):
            ref, bad = refs[s].decode(fr[s], 16)
            assert bad == 0 and np.array_equal(ref[0], out[s]), (k, br, s)


def test_dec_single_frame_api_equals_reference():
    import audio_codec_b200 as ac
    fs, dms, br, nf, N = 32000, 100, 64000, 25, 320
    pcm = make_streams(1, nf * N, fs, N, 16, first_stream=11)
    frames = gpu_encode(fs, dms, br, pcm)[0]
    for bits in (16, 24):
        d = ac.Decoder(fs, 1, dms)
        assert d.frame_len == N and d.delay == CheckDec(fs, 1, dms).delay
        got = np.concatenate([d.dec(frames[f], bits)[0][0] for f in range(nf)])
        ref, bad = CheckDec(fs, 1, dms).decode(frames, bits)
        assert bad == 0 and np.array_equal(ref[0], got)


def test_dec_device_pointers_and_small_chunks():
    import torch
    import audio_codec_b200 as ac
    fs, dms, br, S, nf, N = 48000, 100, 96000, 96, 40, 480
    pcm = make_streams_fast(S, nf * N, fs, N, 16)
    frames = gpu_encode(fs, dms, br, pcm)
    nb = frames.shape[2]
    want, rc = gpu_decode(fs, dms, frames)
    assert rc == 0
    ac.load_library().lc3plus_dec_set_chunk_frames(96 * 7)   # 7 frames per chunk
    try:
        d_in = torch.from_numpy(frames).cuda()
        d_out = torch.zeros((S, nf * N), dtype=torch.int16, device="cuda")
        decs = [ac.Decoder(fs, 1, dms) for _ in range(S)]
        n0 = ac.dec_kernel_launches()
        rc = ac.dec_batch(decs, nf, d_in.data_ptr(), nf * nb, np.full(S, nb, np.int32), d_out.data_ptr(), nf * N * 2, 16)
        assert rc == 0 and ac.dec_kernel_launches() - n0 >= 3 * 6
        assert np.array_equal(d_out.cpu().numpy(), want[:, 0])
    finally:
        ac.load_library().lc3plus_dec_set_chunk_frames(0)


def test_dec_bad_frames_flagged_like_the_reference():
    fs, dms, br, S, nf, N = 32000, 100, 64000, 40, 40, 320
    pcm = make_streams(S, nf * N, fs, N, 16)
    frames = gpu_encode(fs, dms, br, pcm).copy()
    rng = np.random.RandomState(3)
    for s in range(S):
        for f in range(5, nf, 6):
            frames[s, f, rng.randint(0, frames.shape[2], 5)] ^= rng.randint(1, 256, 5).astype(np.uint8)
    bfi = np.zeros((S, nf), dtype=np.uint8)
    out, rc = gpu_decode(fs, dms, frames, bfi=bfi)
    total = 0
    for s in range(S):
        codes = []
        ref, bad = CheckDec(fs, 1, dms).decode(frames[s], 16, codes)
        assert [int(c != 0) for c in codes] == [int(v) for v in bfi[s]], s
        first = codes.index(2) if 2 in codes else nf
        assert np.array_equal(ref[0][:first * N], out[s, 0][:first * N])
        total += bad
    assert total > 0 and rc == 2


def test_dec_lost_stream_and_argument_checks():
    import audio_codec_b200 as ac
    L = ac.load_library()
    fs, dms, N, nf = 16000, 100, 160, 6
    d = ac.Decoder(fs, 1, dms)
    out = np.ones((1, nf * N), dtype=np.int16)
    rc = ac.dec_batch([d], nf, np.zeros(8, np.uint8), 0, np.zeros(1, np.int32), out, nf * N * 2, 16)
    assert rc == 2 and not out.any()
    with pytest.raises(ac.LC3Error) as e:
        ac.dec_batch([d], nf, np.zeros(19 * nf, np.uint8), 19 * nf, np.array([19], np.int32), out, nf * N * 2, 16)
    assert e.value.code == 7
    with pytest.raises(ac.LC3Error) as e:
        ac.dec_batch([d, d], nf, np.zeros(2 * 40 * nf, np.uint8), 40 * nf, np.array([40, 40], np.int32), np.zeros((2, nf * N), np.int16), nf * N * 2, 16)
    assert e.value.code == 1
    assert L.lc3_dec_set_ep_enabled(d.h, 1) == 9 and L.lc3_dec_set_ep_enabled(d.h, 0) == 0
    assert L.lc3_dec_set_frame_dms(d.h, 30) == 10
    assert L.lc3_dec_init(None, 48000, 1, 1) == 3 and L.lc3_dec_init(d.h, 48000, 1, 7) == 8
    assert L.lc3_dec_get_size(48000, 2, 1) > 0 and L.lc3_dec_get_size(48000, 3, 1) == 0


@pytest.mark.parametrize("fs,dms,br,S,nf", [(48000, 100, 128000, 48, 2500), (48000, 100, 64000, 48, 2200), (16000, 50, 32000, 64, 4000),
                                            (32000, 25, 96000, 40, 6000), (44100, 100, 96000, 24, 2100)])
def test_dec_long_runs_equal_reference(fs, dms, br, S, nf):
    """2000+ frames per stream in chunks of three calls (the exponents of the overlap and LTPF memories wander over a long run),
    against the reference decoder run by oracle/_ref/refbench_dec on the same payloads"""
    exe = os.path.join(os.path.dirname(HERE), "oracle", "_ref", "refbench_dec")
    if not os.path.exists(exe):
        pytest.skip("oracle/_ref/refbench_dec not built")
    import subprocess
    import tempfile
    N = (48000 if fs == 44100 else fs) // 100 * dms // 100
    pcm = make_streams_fast(S, nf * N, fs, N, 16, n_unique=S)
    pcm[S // 2:] = make_bursts(S - S // 2, nf * N, fs, 16, seed=21)
    frames = gpu_encode(fs, dms, br, pcm)
    out, rc = gpu_decode(fs, dms, frames, 16, cuts=(nf // 3, nf // 3 * 2 + 1))
    assert rc == 0
    with tempfile.TemporaryDirectory() as d:
        frames.tofile(os.path.join(d, "in.bin"))
        r = subprocess.run([exe, os.path.join(d, "in.bin"), str(fs), str(dms), str(frames.shape[2]), str(S), str(nf), str(min(S, os.cpu_count() or 1)),
                            os.path.join(d, "out.pcm")], capture_output=True, text=True, check=True).stdout.split()
        assert int(r[4]) == 0
        ref = np.fromfile(os.path.join(d, "out.pcm"), dtype=np.int16).reshape(S, nf * N)
    assert np.array_equal(ref, out[:, 0])


def test_dec_reinit_resets_state_and_returns_slots():
    """the reference owns no memory, so callers re-initialise a decoder block without telling anybody: the block's device
    state must start over and its old slots must go back to the pool (5000 re-initialisations stay inside the first pool)"""
    import audio_codec_b200 as ac
    L = ac.load_library()
    fs, dms, br, nf, N = 24000, 100, 48000, 12, 240
    pcm = make_streams(1, nf * N, fs, N, 16, first_stream=4)
    frames = gpu_encode(fs, dms, br, pcm)
    nb = frames.shape[2]
    d = ac.Decoder(fs, 1, dms)
    out1 = np.zeros((1, nf * N), np.int16)
    assert ac.dec_batch([d], nf, frames, nf * nb, np.array([nb], np.int32), out1, nf * N * 2, 16) == 0
    for _ in range(5000):
        assert L.lc3_dec_init(d.h, fs, 1, 1) == 0
    assert L.lc3_dec_set_frame_dms(d.h, dms) == 0
    out2 = np.zeros((1, nf * N), np.int16)
    assert ac.dec_batch([d], nf, frames, nf * nb, np.array([nb], np.int32), out2, nf * N * 2, 16) == 0
    assert np.array_equal(out1, out2)
    out3 = np.zeros((1, nf * N), np.int16)      # without re-initialising, the second pass continues the stream: different PCM
    assert ac.dec_batch([d], nf, frames, nf * nb, np.array([nb], np.int32), out3, nf * N * 2, 16) == 0
    assert not np.array_equal(out1, out3)


def test_dec_concurrent_calls_from_threads():
    """four host threads decode their own streams at the same time (ctypes releases the GIL): the per-device context serialises
    the calls; initialisations of new decoders run alongside"""
    import threading
    import audio_codec_b200 as ac
    fs, dms, br, S, nf, N = 16000, 100, 32000, 24, 30, 160
    pcm = make_streams(4 * S, nf * N, fs, N, 16, silence_frames=2)
    frames = gpu_encode(fs, dms, br, pcm)
    nb = frames.shape[2]
    outs, errs = [None] * 4, []

    def work(t):
        try:
            for rep in range(3):
                decs = [ac.Decoder(fs, 1, dms) for _ in range(S)]
                o = np.zeros((S, nf * N), dtype=np.int16)
                fr = np.ascontiguousarray(frames[t * S:(t + 1) * S])
                assert ac.dec_batch(decs, nf, fr, nf * nb, np.full(S, nb, np.int32), o, nf * N * 2, 16) == 0
                outs[t] = o
                for d in decs:
                    d.close()
        except Exception as e:   # noqa: BLE001
            errs.append(e)
    th = [threading.Thread(target=work, args=(t,)) for t in range(4)]
    for t in th:
        t.start()
    for t in th:
        t.join()
    assert not errs, errs
    for t in range(4):
        for s in (0, S // 2, S - 1):
            ref, bad = CheckDec(fs, 1, dms).decode(frames[t * S + s], 16)
            assert bad == 0 and np.array_equal(ref[0], outs[t][s]), (t, s)


@pytest.mark.parametrize("item,fs,br", [("thetest8", 8000, 32000), ("thetest16", 16000, 32000), ("thetest24", 24000, 48000),
                                        ("thetest32", 32000, 48000), ("thetest44", 44100, 64000), ("thetest48", 48000, 64000)])
def test_dec_etsi_items_equal_reference(item, fs, br):
    """the ETSI package's own test items (tests/golden/etsi_full.npz: speech and music, about 10 s each) at the package's bitrates
    (ETSI/testvec/testvecCheck.pl), no frame erasures: real pitch contours drive every LTPF transition"""
    z = np.load(os.path.join(HERE, "golden", "etsi_full.npz"))
    N = (48000 if fs == 44100 else fs) // 100
    x = z[item]
    nf = len(x) // N
    pcm = np.ascontiguousarray(x[:nf * N][None, :])
    frames = gpu_encode(fs, 100, br, pcm)
    for bits in (16, 24):
        out, rc = gpu_decode(fs, 100, frames, bits, cuts=(nf // 2,))
        ref, bad = CheckDec(fs, 1, 100).decode(frames[0], bits)
        assert rc == 0 and bad == 0 and np.array_equal(ref[0], out[0, 0]), bits


def test_dec_full_size_round_trip_c1():
    """BASELINE.json config c1 geometry (4096 streams x 48 kHz / 10 ms / 128 kbps), 100 frames: encode -> decode on the device,
    a sample of streams against the reference decoder, every stream against the delayed input (a codec at 128 kbps keeps the
    signal: SNR well above 15 dB on the harmonic test streams)."""
    import torch
    import audio_codec_b200 as ac
    fs, dms, br, S, nf, N = 48000, 100, 128000, 4096, 100, 480
    pcm = make_streams_fast(S, nf * N, fs, N, 16)
    frames = gpu_encode(fs, dms, br, pcm)
    nb = frames.shape[2]
    decs = ac.DecoderArray([ac.Decoder(fs, 1, dms) for _ in range(S)])
    d_in = torch.from_numpy(frames).cuda()
    d_out = torch.zeros((S, nf * N), dtype=torch.int16, device="cuda")
    rc = ac.dec_batch(decs, nf, d_in.data_ptr(), nf * nb, np.full(S, nb, np.int32), d_out.data_ptr(), nf * N * 2, 16)
    assert rc == 0
    out = d_out.cpu().numpy()
    if True:
        for s in (0, 7, 63, 64, 1000, 4095):
            ref, bad = CheckDec(fs, 1, dms).decode(frames[s], 16)
            assert bad == 0 and np.array_equal(ref[0], out[s]), s
    delay = decs.decoders[0].delay
    a = pcm[:, 10 * N:(nf - 1) * N].astype(np.float64)
    b = out[:, 10 * N + delay:(nf - 1) * N + delay].astype(np.float64)
    snr = 10 * np.log10((a ** 2).sum(1) / np.maximum(((a - b) ** 2).sum(1), 1e-9))
    assert snr.min() > 15, snr.min()
