"""Parity tests proper: the CUDA path, through the C ABI, against the oracle on the same seeded inputs.
Bar: every byte of every frame of every stream equal (integer/byte work => bit-exact)."""
import numpy as np
import pytest

import lc3ref
from audio_codec_b200.synth import make_streams, make_streams_fast, make_torture

pytestmark = pytest.mark.gpu


def _frame_len(fs, dms):
    return (48000 if fs == 44100 else fs) // 100 * dms // 100


def _oracle(fs, ch, dms, br, bits, pcm, bfi=None):
    return lc3ref.OracleEnc(fs, ch, dms, br).encode(pcm, bits, bfi)


def _batch(ac, encs, nf, pcm, bits, num_bytes=None):
    nb = max(e.num_bytes for e in encs)
    out = np.zeros((len(encs), nf, nb), dtype=np.uint8)
    ac.enc_batch(encs, nf, pcm, pcm[0].nbytes, bits, out, nf * nb, num_bytes)
    return out


CASES = [(48000, 1, 100, 128000, 16), (16000, 1, 50, 32000, 16), (48000, 2, 25, 256000, 16),
         (8000, 1, 100, 16000, 16), (24000, 2, 50, 96000, 24), (32000, 1, 100, 80000, 16),
         (44100, 1, 100, 96000, 16), (48000, 1, 100, 128000, 24), (32000, 1, 25, 96000, 16),
         (48000, 1, 100, 400000 * 8 // 10, 16), (48000, 1, 100, 16000, 16), (16000, 1, 100, 16000, 16)]


@pytest.mark.parametrize("fs,ch,dms,br,bits", CASES)
def test_batch_equals_oracle(fs, ch, dms, br, bits):
    import audio_codec_b200 as ac
    N, nf, S = _frame_len(fs, dms), 24, 40   # 40 streams: one full tile and a ragged one
    pcm = np.stack([make_streams(ch, nf * N, fs, N, bits, first_stream=s * ch, silence_frames=8) for s in range(S)])
    encs = [ac.Encoder(fs, ch, dms, br) for _ in range(S)]
    out = _batch(ac, encs, nf, pcm, bits)
    for s in range(S):
        want = _oracle(fs, ch, dms, br, bits, pcm[s])
        assert np.array_equal(out[s, :, :want.shape[1]], want), f"stream {s}"
    for e in encs:
        e.close()


def test_state_carries_across_calls_and_chunks():
    """Many small calls, one big call and a big call cut into internal chunks give the same bytes."""
    import audio_codec_b200 as ac
    fs, dms, br, N, nf, S = 48000, 100, 128000, 480, 20, 5
    pcm = make_streams(S, nf * N, fs, N)[:, None, :]
    want = np.stack([_oracle(fs, 1, dms, br, 16, pcm[s]) for s in range(S)])
    e1 = [ac.Encoder(fs, 1, dms, br) for _ in range(S)]
    assert np.array_equal(_batch(ac, e1, nf, pcm, 16), want)
    e2 = [ac.Encoder(fs, 1, dms, br) for _ in range(S)]
    parts = [_batch(ac, e2, 4, np.ascontiguousarray(pcm[:, :, k * 4 * N:(k + 1) * 4 * N]), 16) for k in range(5)]
    assert np.array_equal(np.concatenate(parts, axis=1), want)
    ac.set_chunk_frames(32 * 3)   # 3 frames per internal chunk
    e3 = [ac.Encoder(fs, 1, dms, br) for _ in range(S)]
    got = _batch(ac, e3, nf, pcm, 16)
    ac.set_chunk_frames(1 << 19)
    assert np.array_equal(got, want)


def test_single_frame_api_equals_oracle():
    """lc3_enc16 / lc3_enc24 one frame at a time, stereo, with a 'lost frame' marker (fx/lc3.c:253)."""
    import audio_codec_b200 as ac
    for bits in (16, 24):
        fs, ch, dms, br, nf = 32000, 2, 100, 128000, 12
        N = _frame_len(fs, dms)
        pcm = make_streams(ch, nf * N, fs, N, bits, first_stream=4)
        bfi = [0, 0, 0, 1, 0, 0, 0, 0, 1, 0, 0, 0]
        want = _oracle(fs, ch, dms, br, bits, pcm, bfi)
        e = ac.Encoder(fs, ch, dms, br)
        got = np.stack([e.enc(pcm[:, f * N:(f + 1) * N], bits, lost=bool(bfi[f])) for f in range(nf)])
        assert np.array_equal(got, want)


def test_bitrate_switching_between_calls():
    import audio_codec_b200 as ac
    fs, dms, N = 48000, 100, 480
    sched = [128000, 32000, 320000, 96000, 64000, 128000]
    pcm = make_streams(3, len(sched) * 6 * N, fs, N, first_stream=1)
    encs = [ac.Encoder(fs, 1, dms) for _ in range(3)]
    oras = [lc3ref.OracleEnc(fs, 1, dms) for _ in range(3)]
    for k, br in enumerate(sched):
        seg = np.ascontiguousarray(pcm[:, None, k * 6 * N:(k + 1) * 6 * N])
        for e, o in zip(encs, oras):
            assert e.set_bitrate(br) == 0 and o.set_bitrate(br) == 0
        got = _batch(ac, encs, 6, seg, 16)
        for s in range(3):
            want = oras[s].encode(seg[s])
            assert np.array_equal(got[s, :, :want.shape[1]], want), (k, s)


def test_mixed_configuration_batch():
    """BASELINE config 5: one call with streams of different fs / frame size / bitrate / channel count."""
    import audio_codec_b200 as ac
    cfgs = [(8000, 1, 25, 64000), (24000, 1, 50, 48000), (32000, 2, 100, 160000), (44100, 1, 100, 96000),
            (48000, 1, 25, 192000), (48000, 1, 100, 128000), (16000, 1, 100, 32000), (48000, 2, 50, 128000)] * 3
    dur_ms = 100
    encs, pcms, sizes = [], [], []
    for i, (fs, ch, dms, br) in enumerate(cfgs):
        encs.append(ac.Encoder(fs, ch, dms, br))
        N = _frame_len(fs, dms)
        nf = dur_ms * 10 // dms
        pcms.append(make_streams(ch, nf * N, fs, N, 16, first_stream=i))
        sizes.append(nf)
    # lc3plus_enc_batch takes one n_frames per call: group by frame duration, mixed fs/bitrate/channels inside
    for dms in (25, 50, 100):
        idx = [i for i, c in enumerate(cfgs) if c[2] == dms]
        nf = dur_ms * 10 // dms
        stride = max(pcms[i].nbytes for i in idx)
        buf = np.zeros((len(idx), stride // 2), dtype=np.int16)
        for k, i in enumerate(idx):
            buf[k, :pcms[i].size] = pcms[i].reshape(-1)
        ostride = nf * max(encs[i].num_bytes for i in idx)
        out = np.zeros((len(idx), ostride), dtype=np.uint8)
        nbytes = np.zeros(len(idx) * nf, dtype=np.int32)
        ac.enc_batch([encs[i] for i in idx], nf, buf, stride, 16, out, ostride, nbytes)
        for k, i in enumerate(idx):
            fs, ch, d, br = cfgs[i]
            want = _oracle(fs, ch, d, br, 16, pcms[i])
            assert np.array_equal(out[k, :want.size].reshape(want.shape), want), cfgs[i]
            assert (nbytes[k * nf:(k + 1) * nf] == want.shape[1]).all()


def test_device_pointers_and_large_batch_properties():
    """Full-size config: 4096 streams x 48 kHz/10 ms/128 kbps with device-resident PCM and output.
    Size-independent checks: a sample of streams equals the oracle; streams with identical PCM give
    identical bytes; re-encoding from fresh encoders is deterministic."""
    import torch
    import audio_codec_b200 as ac
    fs, dms, br, N, nf, S = 48000, 100, 128000, 480, 12, 4096
    pcm = make_streams_fast(S, nf * N, fs, N, n_unique=64)
    pcm[64:128] = pcm[0:64]            # duplicates: must encode identically
    d_pcm = torch.from_numpy(pcm).cuda()
    nb = br * N // (8 * fs)
    outs = []
    for _ in range(2):
        encs = [ac.Encoder(fs, 1, dms, br) for _ in range(S)]
        d_out = torch.zeros((S, nf, nb), dtype=torch.uint8, device="cuda")
        ac.enc_batch(encs, nf, d_pcm.data_ptr(), nf * N * 2, 16, d_out.data_ptr(), nf * nb)
        torch.cuda.synchronize()
        outs.append(d_out.cpu().numpy())
        for e in encs:
            e.close()
    assert np.array_equal(outs[0], outs[1])
    assert np.array_equal(outs[0][64:128], outs[0][0:64])
    for s in (0, 7, 31, 32, 63, 1000, 4095):
        assert np.array_equal(outs[0][s], _oracle(fs, 1, dms, br, 16, pcm[s:s + 1])), s


def test_bandwidth_control():
    import audio_codec_b200 as ac
    fs, dms, br, N, nf = 48000, 100, 96000, 480, 10
    pcm = make_streams(1, nf * N, fs, N, first_stream=5)
    e = ac.Encoder(fs, 1, dms, br)
    o = lc3ref.OracleEnc(fs, 1, dms, br)
    assert e.set_bandwidth(8000) == 0 and o.set_bandwidth(8000) == 0
    got = _batch(ac, [e], nf, pcm[None], 16)[0]
    assert np.array_equal(got, o.encode(pcm))


def test_error_codes():
    import audio_codec_b200 as ac
    e = ac.Encoder(16000, 1, 100)
    with pytest.raises(ac.LC3Error) as ex:
        e.enc(np.zeros((1, 160), dtype=np.int16))
    assert ex.value.code == 14                        # LC3_BITRATE_UNSET_ERROR (fx/lc3.c:252)
    assert e.set_bitrate(15999, check=False) == 6     # below 20 bytes/frame
    assert e.set_bitrate(320001, check=False) == 6
    assert e.set_bitrate(32000) == 0
    assert e.set_frame_dms(50) == 15                  # LC3_BITRATE_SET_ERROR (fx/lc3.c:219)
    assert e.num_bytes == 40 and e.frame_len == 160 and e.delay == 160 - 2 * 60


@pytest.mark.parametrize("fs,dms,br,bits", [(48000, 100, 128000, 16), (48000, 100, 320000, 24), (16000, 50, 32000, 16),
                                            (48000, 25, 256000, 16), (32000, 100, 64000, 16), (8000, 100, 16000, 24),
                                            (24000, 25, 96000, 24), (44100, 100, 160000, 16)])
def test_adversarial_signals(fs, dms, br, bits):
    """Rail-to-rail noise, squares, DC, impulses, steps: the inputs that reach the saturating operators."""
    import audio_codec_b200 as ac
    N, nf = _frame_len(fs, dms), 20
    pcm = make_torture(nf * N, fs, bits)[:, None, :]
    encs = [ac.Encoder(fs, 1, dms, br) for _ in range(pcm.shape[0])]
    out = _batch(ac, encs, nf, pcm, bits)
    for s in range(pcm.shape[0]):
        want = _oracle(fs, 1, dms, br, bits, pcm[s])
        assert np.array_equal(out[s, :, :want.shape[1]], want), f"stream {s}"


@pytest.mark.parametrize("fs,dms,br,bits", [(48000, 100, 128000, 16), (32000, 100, 32000, 16), (48000, 50, 96000, 24), (16000, 25, 64000, 16)])
def test_transient_rich_signals(fs, dms, br, bits):
    """Gated tones, noise bursts and clicks (synth.make_bursts): TNS, attack handling and pitch changes in most frames."""
    import audio_codec_b200 as ac
    from audio_codec_b200.synth import make_bursts
    N, nf = _frame_len(fs, dms), 40
    pcm = make_bursts(40, nf * N, fs, bits)[:, None, :]
    encs = [ac.Encoder(fs, 1, dms, br) for _ in range(pcm.shape[0])]
    out = _batch(ac, encs, nf, pcm, bits)
    for s in range(pcm.shape[0]):
        want = _oracle(fs, 1, dms, br, bits, pcm[s])
        assert np.array_equal(out[s, :, :want.shape[1]], want), f"stream {s}"


def test_error_protection_mixed_batch():
    """One batch call whose streams use different error-protection modes (and none), mono and stereo, jointly coded and
    one slot per channel, with partial concealment (slots >= 80 bytes, modes 3/4): every stream equals the oracle
    (fx/al_fec.c fec_encoder, fx/reorder_bitstream_fx.c, fx/enc_lc3.c:247-312)."""
    import audio_codec_b200 as ac
    fs, dms, nf = 48000, 100, 12
    N = _frame_len(fs, dms)
    cfgs = [(1, 128000, 0, 0), (1, 128000, 4, 0), (1, 128000, 3, 2), (1, 64000, 4, 1), (1, 48000, 2, 3), (1, 32000, 1, 0),
            (2, 96000, 3, 0), (2, 128000, 4, 2), (2, 256000, 4, 1), (1, 240000, 3, 3), (1, 240000, 4, 0), (2, 192000, 2, 0)] * 3
    encs, oras, pcms = [], [], []
    for i, (ch, br, ep, epmr) in enumerate(cfgs):
        e, o = ac.Encoder(fs, ch, dms), lc3ref.OracleEnc(fs, ch, dms)
        assert e.set_ep_mode(ep) == o.set_ep_mode(ep) == 0
        assert e.set_bitrate(br) == o.set_bitrate(br) == 0
        assert e.set_ep_mode_request(epmr) == o.set_ep_mode_request(epmr) == 0
        assert e.num_bytes == o.num_bytes
        encs.append(e); oras.append(o)
        pcms.append(make_streams(ch, nf * N, fs, N, 16, first_stream=i, silence_frames=2))
    stride = 2 * nf * N * 2
    buf = np.zeros((len(cfgs), stride // 2), dtype=np.int16)
    for k, p in enumerate(pcms):
        buf[k, :p.size] = p.reshape(-1)
    ostride = nf * max(e.num_bytes for e in encs)
    out = np.zeros((len(cfgs), ostride), dtype=np.uint8)
    ac.enc_batch(encs, nf, buf, stride, 16, out, ostride)
    for k, o in enumerate(oras):
        want = o.encode(pcms[k])
        assert np.array_equal(out[k, :want.size].reshape(want.shape), want), cfgs[k]


def test_error_protection_switching_and_codes():
    """lc3_enc_set_ep_mode between frames (the -epf switching of fx/codec_exe.c:351-363) and its error returns."""
    import audio_codec_b200 as ac
    fs, N = 32000, 320
    e, o = ac.Encoder(fs, 1, 100, 96000), lc3ref.OracleEnc(fs, 1, 100, 96000)
    pcm = make_streams(1, 30 * N, fs, N, 16, first_stream=4)
    for f in range(30):
        ep = (0, 4, 3, 1, 2, 4)[(f // 2) % 6]
        assert e.set_ep_mode(ep) == o.set_ep_mode(ep) == 0
        got = e.enc(pcm[:, f * N:(f + 1) * N])
        assert np.array_equal(got, o.encode(pcm[:, f * N:(f + 1) * N])[0]), (f, ep)
    assert e.set_ep_mode(5) == 9 and e.set_ep_mode_request(4) == 12     # LC3_EPMODE_ERROR, LC3_EPMR_ERROR
    e2 = ac.Encoder(16000, 1, 100)
    assert e2.set_ep_mode(3) == 0
    assert e2.set_bitrate(16000, check=False) == 6                     # 20-byte slot is below FEC_SLOT_BYTES_MIN
    assert e2.set_bitrate(32000) == 0 and e2.num_bytes == 40
