#!/usr/bin/env python
"""One batch with error protection on (k_fec runs after k_back): python tools/fec_run.py [epmode] -- for ncu captures of k_fec."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import audio_codec_b200 as ac  # noqa: E402
from audio_codec_b200.synth import make_streams_fast  # noqa: E402

ep = int(sys.argv[1]) if len(sys.argv) > 1 else 4
S, F, fs, dms, br, N = 4096, 100, 48000, 100, 128000, 480
encs = []
for _ in range(S):
    e = ac.Encoder(fs, 1, dms)
    assert e.set_ep_mode(ep) == 0
    e.set_bitrate(br)
    encs.append(e)
encs = ac.EncoderArray(encs)
nb = encs.encoders[0].num_bytes
pcm = torch.from_numpy(make_streams_fast(S, F * N, fs, N)).cuda()
out = torch.zeros((S, F, nb), dtype=torch.uint8, device="cuda")
ac.set_chunk_frames(S * F)
for _ in range(2):
    ac.enc_batch(encs, F, pcm.data_ptr(), F * N * 2, 16, out.data_ptr(), F * nb)
torch.cuda.synchronize()
print("ok", nb)
