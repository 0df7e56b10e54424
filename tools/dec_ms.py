#!/usr/bin/env python3
"""Decoder timing on the GPU: tools/dec_ms.py [streams] [frames] [fs] [dms] [bitrate].  Payloads come from the GPU encoder; the
decode is timed device-resident (CUDA events inside the library, per pass) and end to end with host buffers (wall clock around
the synchronous call); a sample of streams is checked against the compiled reference decoder when it is present."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import audio_codec_b200 as ac  # noqa: E402
from audio_codec_b200.synth import make_streams_fast  # noqa: E402


def main():
    a = [int(x) for x in sys.argv[1:]]
    S, nf, fs, dms, br = (a + [4096, 1000, 48000, 100, 128000][len(a):])[:5]
    N = (48000 if fs == 44100 else fs) // 100 * dms // 100
    pcm = make_streams_fast(S, nf * N, fs, N, 16)
    encs = ac.EncoderArray([ac.Encoder(fs, 1, dms, br) for _ in range(S)])
    nb = encs.encoders[0].num_bytes
    d_pcm = torch.from_numpy(pcm).cuda()
    d_bytes = torch.zeros((S, nf * nb), dtype=torch.uint8, device="cuda")
    ac.enc_batch(encs, nf, d_pcm.data_ptr(), nf * N * 2, 16, d_bytes.data_ptr(), nf * nb)
    del d_pcm
    d_out = torch.zeros((S, nf * N), dtype=torch.int16, device="cuda")
    nbs = np.full(S, nb, np.int32)
    res = {}
    for rep in range(3):
        decs = ac.DecoderArray([ac.Decoder(fs, 1, dms) for _ in range(S)])
        ac.dec_profile_enable(True)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        rc = ac.dec_batch(decs, nf, d_bytes.data_ptr(), nf * nb, nbs, d_out.data_ptr(), nf * N * 2, 16)
        dt = time.perf_counter() - t0
        p = ac.dec_profile_read()
        res = {"rc": rc, "wall_ms": dt * 1e3, "frames_per_s": S * nf / dt, "first_chunk_ms": p}
        for d in decs.decoders:
            d.close()
    # end to end: payloads and PCM in PINNED host memory, copies inside the timed call
    t_in = torch.empty((S, nf * nb), dtype=torch.uint8, pin_memory=True)
    t_in.copy_(d_bytes)
    t_out = torch.empty((S, nf * N), dtype=torch.int16, pin_memory=True)
    h_bytes, h_out = t_in.numpy(), t_out.numpy()
    for rep in range(2):
        decs = ac.DecoderArray([ac.Decoder(fs, 1, dms) for _ in range(S)])
        t0 = time.perf_counter()
        ac.dec_batch(decs, nf, h_bytes, nf * nb, nbs, h_out, nf * N * 2, 16)
        dt = time.perf_counter() - t0
        res["e2e_frames_per_s"] = S * nf / dt
        res["e2e_pcie_GBps"] = S * nf * (nb + N * 2) / dt / 1e9
        for d in decs.decoders:
            d.close()
    assert np.array_equal(h_out, d_out.cpu().numpy())
    try:
        import lc3ref
        if lc3ref.have_ref():
            fr = h_bytes.reshape(S, nf, nb)
            t0 = time.perf_counter()
            for s in (0, S // 2, S - 1):
                ref, bad = lc3ref.RefDec(fs, 1, dms).decode(fr[s][:200], 16)
                assert bad == 0 and np.array_equal(ref[0], h_out[s][:200 * N]), s
            res["verified_streams"] = 3
            res["ref_dec_frames_per_s_one_core"] = 3 * min(nf, 200) / (time.perf_counter() - t0)
    except ImportError:
        pass
    res["config"] = {"streams": S, "frames": nf, "fs": fs, "dms": dms, "bitrate": br, "nbytes": nb}
    print(json.dumps(res))


if __name__ == "__main__":
    main()
