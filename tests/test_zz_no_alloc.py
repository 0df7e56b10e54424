"""lc3plus_enc_fresh_noalloc (include/lc3plus_b200.h): the allocation-free entry point, through the C ABI, against the
existing entry point lc3plus_enc_batch on fresh encoders.  Integer/byte work: the bar is bit-exact.  Two calls in a row
share the static device scratch, so the second call (other signal, other configuration) must not see the first one's
state.  Collected last on purpose (file name)."""
import numpy as np
import pytest

from audio_codec_b200.synth import make_streams

pytestmark = pytest.mark.gpu


def _frame_len(fs, dms):
    return (48000 if fs == 44100 else fs) // 100 * dms // 100


def _reference_entry_point(ac, fs, dms, br, bits, pcm, nf):
    encs = [ac.Encoder(fs, 1, dms, br) for _ in range(pcm.shape[0])]
    nb = encs[0].num_bytes
    out = np.zeros((pcm.shape[0], nf, nb), dtype=np.uint8)
    ac.enc_batch(encs, nf, pcm, pcm[0].nbytes, bits, out, nf * nb)
    for e in encs:
        e.close()
    return out


def test_noalloc_equals_enc_batch_two_calls_in_a_row():
    import torch
    import audio_codec_b200 as ac
    calls = [(48000, 100, 128000, 16, 40, 16, 0), (48000, 100, 128000, 16, 40, 16, 100),   # same shape, other signal
             (16000, 50, 32000, 16, 64, 16, 3), (48000, 100, 96000, 24, 32, 32, 7), (32000, 25, 96000, 16, 33, 12, 11)]
    for k, (fs, dms, br, bits, S, nf, first) in enumerate(calls):
        N = _frame_len(fs, dms)
        pcm = make_streams(S, nf * N, fs, N, bits, first_stream=first, silence_frames=4)
        want = _reference_entry_point(ac, fs, dms, br, bits, pcm, nf)
        nb = want.shape[2]
        d_pcm = torch.from_numpy(pcm).cuda()
        d_out = torch.zeros((S, nf, nb), dtype=torch.uint8, device="cuda")
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        ac.enc_fresh_noalloc(fs, dms, br, bits, S, nf, d_pcm.data_ptr(), pcm[0].nbytes, d_out.data_ptr(), nf * nb)
        torch.cuda.synchronize()
        # (the very first call makes the driver load the set-up kernel's code: lazy module loading, not an allocation of ours)
        assert k == 0 or torch.cuda.mem_get_info()[0] == free0, "the call changed the amount of free device memory"
        got = d_out.cpu().numpy()
        assert np.array_equal(got, want), (fs, dms, br, bits, S, nf)


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    import torch
    import audio_codec_b200 as ac
    d = torch.zeros(1 << 20, dtype=torch.uint8, device="cuda")
    for S, nf in ((65, 1), (64, 17), (33, 17), (1, 33)):
        with pytest.raises(ac.LC3Error) as ex:
            ac.enc_fresh_noalloc(48000, 100, 128000, 16, S, nf, d.data_ptr(), nf * 960, d.data_ptr(), nf * 160)
        assert ex.value.code == 1 and "static scratch" in str(ex.value)
    host = np.zeros(960 * 4, dtype=np.int16)
    with pytest.raises(ac.LC3Error) as ex:
        ac.enc_fresh_noalloc(48000, 100, 128000, 16, 1, 1, host.ctypes.data, 960, d.data_ptr(), 160)
    assert "device pointers" in str(ex.value)
    with pytest.raises(ac.LC3Error) as ex:
        ac.enc_fresh_noalloc(48000, 100, 1000, 16, 1, 1, d.data_ptr(), 960, d.data_ptr(), 160)
    assert ex.value.code == 6   # LC3_BITRATE_ERROR, as lc3_enc_set_bitrate
