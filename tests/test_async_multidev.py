"""The host layer beyond one synchronous call on one GPU: lc3plus_enc_batch_async / lc3plus_sync (consecutive calls form one
pipeline, state carried from call to call without draining the device), lc3plus_set_device (one context per device, a batch
whose encoders live on several devices), and the handle registry (re-initialising an encoder block leaks nothing)."""
import numpy as np
import pytest

import lc3ref
from audio_codec_b200.synth import make_streams

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("device_buffers", [False, True])
def test_async_calls_pipeline_and_carry_state(device_buffers):
    import torch
    import audio_codec_b200 as ac
    fs, dms, br, N, S, seg, nseg = 48000, 100, 128000, 480, 37, 9, 5
    pcm = make_streams(S, seg * nseg * N, fs, N, first_stream=2, silence_frames=12)
    encs = ac.EncoderArray([ac.Encoder(fs, 1, dms, br) for _ in range(S)])
    nb = encs.encoders[0].num_bytes
    ac.set_chunk_frames(64 * 2)     # several chunks per call, so that calls interleave chunk by chunk
    try:
        segs = [np.ascontiguousarray(pcm[:, k * seg * N:(k + 1) * seg * N]) for k in range(nseg)]
        if device_buffers:
            ins = [torch.from_numpy(x).cuda() for x in segs]
            outs = [torch.zeros((S, seg, nb), dtype=torch.uint8, device="cuda") for _ in range(nseg)]
            for x, o in zip(ins, outs):
                ac.enc_batch(encs, seg, x.data_ptr(), seg * N * 2, 16, o.data_ptr(), seg * nb, wait=False)
            ac.sync()
            got = np.concatenate([o.cpu().numpy() for o in outs], axis=1)
        else:
            ins = [torch.from_numpy(x).pin_memory() for x in segs]
            outs = [torch.zeros((S, seg, nb), dtype=torch.uint8).pin_memory() for _ in range(nseg)]
            for x, o in zip(ins, outs):
                ac.enc_batch(encs, seg, x.data_ptr(), seg * N * 2, 16, o.data_ptr(), seg * nb, wait=False)
            ac.sync()
            got = np.concatenate([o.numpy() for o in outs], axis=1)
    finally:
        ac.set_chunk_frames(0)
    for s in range(S):
        want = lc3ref.OracleEnc(fs, 1, dms, br).encode(pcm[s:s + 1])
        assert np.array_equal(got[s], want), s


def test_async_then_bitrate_switch_then_sync_call():
    """A configuration change between asynchronous calls must not reach frames already queued."""
    import audio_codec_b200 as ac
    fs, dms, N, S, seg = 48000, 100, 480, 6, 8
    pcm = make_streams(S, 3 * seg * N, fs, N, first_stream=9)
    encs = [ac.Encoder(fs, 1, dms, 128000) for _ in range(S)]
    oras = [lc3ref.OracleEnc(fs, 1, dms, 128000) for _ in range(S)]
    outs, want = [], []
    for k, br in enumerate((128000, 32000, 96000)):
        x = np.ascontiguousarray(pcm[:, k * seg * N:(k + 1) * seg * N])
        for e, o in zip(encs, oras):
            assert e.set_bitrate(br) == 0 and o.set_bitrate(br) == 0
        out = np.zeros((S, seg * 160), dtype=np.uint8)    # frame f of a stream at f * num_bytes
        ac.enc_batch(encs, seg, x, seg * N * 2, 16, out, seg * 160, wait=(k == 2))
        outs.append((out, x))
        want.append([o.encode(x[s:s + 1]) for s, o in enumerate(oras)])
    ac.sync()
    for (out, _), w in zip(outs, want):
        for s in range(S):
            assert np.array_equal(out[s, :w[s].size].reshape(w[s].shape), w[s])


def test_reinitialising_a_block_reuses_its_state_slots():
    import ctypes as C
    import torch
    import audio_codec_b200 as ac
    L = ac.load_library()
    e = ac.Encoder(48000, 2, 100, 128000)
    x = make_streams(2, 480, 48000, 480)
    e.enc(x)
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for _ in range(20000):      # the initial pool holds 4096 slots: a leak of two slots per init would grow it three times
        assert L.lc3_enc_init(e.h, 48000, 2) == 0
    assert L.lc3_enc_set_frame_dms(e.h, 100) == 0 and L.lc3_enc_set_bitrate(e.h, 128000) == 0
    got = e.enc(x)
    assert torch.cuda.mem_get_info()[0] == free0
    assert np.array_equal(got, lc3ref.OracleEnc(48000, 2, 100, 128000).encode(x)[0])
    n = C.c_int(0)
    out = np.zeros(400, dtype=np.uint8)
    ptrs = (C.c_void_p * 2)(x[0].ctypes.data, x[1].ctypes.data)
    assert L.lc3_enc16(e.h, ptrs, out.ctypes.data, C.byref(n), None) == 3    # scratch == NULL: LC3_NULL_ERROR (fx/lc3.c:250)


def test_encoders_on_every_device_in_one_batch():
    """lc3plus_set_device + one lc3plus_enc_batch over host buffers whose encoders are spread over all visible GPUs
    (one GPU: the same path with a single context)."""
    import torch
    import audio_codec_b200 as ac
    nd = torch.cuda.device_count()
    fs, dms, br, N, S, nf = 32000, 100, 64000, 320, 24 * nd, 10
    pcm = make_streams(S, nf * N, fs, N, first_stream=5)
    encs = []
    for s in range(S):
        ac.set_device(s * nd // S)          # contiguous shards, like audio_codec_b200.shard.shard_range
        encs.append(ac.Encoder(fs, 1, dms, br))
        assert encs[-1].device == s * nd // S
    ac.set_device(-1)
    nb = encs[0].num_bytes
    out = np.zeros((S, nf, nb), dtype=np.uint8)
    ac.enc_batch(encs, nf, pcm, nf * N * 2, 16, out, nf * nb)
    for s in range(S):
        assert np.array_equal(out[s], lc3ref.OracleEnc(fs, 1, dms, br).encode(pcm[s:s + 1])), s
    # interleaved placement (rows of one device are not consecutive): whole-buffer copies
    encs2 = []
    for s in range(S):
        ac.set_device(s % nd)
        encs2.append(ac.Encoder(fs, 1, dms, br))
    ac.set_device(-1)
    out2 = np.zeros_like(out)
    ac.enc_batch(encs2, nf, pcm, nf * N * 2, 16, out2, nf * nb)
    assert np.array_equal(out2, out)


def test_set_bitrate_many_and_device_profile():
    """lc3plus_enc_set_bitrate_many == lc3_enc_set_bitrate on each encoder (first error returned, the rest applied);
    lc3plus_profile_read_device reports the span of asynchronous calls on the device they ran on."""
    import audio_codec_b200 as ac
    fs, dms, N, S, nf = 48000, 100, 480, 5, 6
    pcm = make_streams(S, 2 * nf * N, fs, N, first_stream=3)
    encs = ac.EncoderArray([ac.Encoder(fs, 1, dms, 64000) for _ in range(S)])
    oras = [lc3ref.OracleEnc(fs, 1, dms, 64000) for _ in range(S)]
    rates = np.array([128000, 32000, 96000, 16000, 320000], dtype=np.int32)
    ac.profile_enable(1)
    outs = []
    for k in range(2):
        x = np.ascontiguousarray(pcm[:, k * nf * N:(k + 1) * nf * N])
        if k == 1:
            ac.set_bitrate_many(encs, rates)
            for o, r in zip(oras, rates):
                assert o.set_bitrate(int(r)) == 0
        out = np.zeros((S, nf * 400), dtype=np.uint8)
        ac.enc_batch(encs, nf, x, nf * N * 2, 16, out, nf * 400, wait=False)
        outs.append((out, [o.encode(x[s:s + 1]) for s, o in enumerate(oras)]))
    ac.sync()
    prof = ac.profile_read_device(encs.encoders[0].device)
    ac.profile_enable(0)
    assert prof["span"] > 0 and prof["calls"] == 2
    for out, want in outs:
        for s in range(S):
            assert np.array_equal(out[s, :want[s].size].reshape(want[s].shape), want[s]), s
    bad = rates.copy()
    bad[2] = 1000                                  # illegal: LC3_BITRATE_ERROR for that one, the others still applied
    with pytest.raises(ac.LC3Error) as ex:
        ac.set_bitrate_many(encs, bad)
    assert ex.value.code == 6
    assert [e.num_bytes for e in encs.encoders] == [160, 40, 120, 20, 400]
