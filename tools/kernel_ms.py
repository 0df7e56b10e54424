#!/usr/bin/env python
"""Per-kernel device milliseconds of one unpipelined step (one chunk = the whole batch, so no two kernels overlap):

    python tools/kernel_ms.py [streams] [frames] [fs] [dms] [bitrate]

Uses the library's own CUDA-event profile (lc3plus_profile_enable / _read).  Prints one JSON line."""
import json
import os
import sys

os.environ.setdefault("LC3PLUS_B200_CHAIN", "1")  # kernels of the chunk strictly one after the other

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import audio_codec_b200 as ac  # noqa: E402
from audio_codec_b200.synth import make_streams  # noqa: E402

S = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
F = int(sys.argv[2]) if len(sys.argv) > 2 else 100
fs = int(sys.argv[3]) if len(sys.argv) > 3 else 48000
dms = int(sys.argv[4]) if len(sys.argv) > 4 else 100
br = int(sys.argv[5]) if len(sys.argv) > 5 else 128000
encs = [ac.Encoder(fs, 1, dms, br) for _ in range(S)]
N, nb = encs[0].frame_len, encs[0].num_bytes
pcm = torch.from_numpy(make_streams(S, F * N, fs, N)).cuda()
out = torch.zeros((S, F, nb), dtype=torch.uint8, device="cuda")
ac.set_chunk_frames(S * F)
for _ in range(3):
    ac.enc_batch(encs, F, pcm.data_ptr(), F * N * 2, 16, out.data_ptr(), F * nb)
torch.cuda.synchronize()
reps = 10
ac.profile_enable(True)
for _ in range(reps):
    ac.enc_batch(encs, F, pcm.data_ptr(), F * N * 2, 16, out.data_ptr(), F * nb)
torch.cuda.synchronize()
p = ac.profile_read()
ac.profile_enable(False)
ks = ("stage", "fir", "scalar", "pitch", "mdct", "sns", "tns", "rate", "quant", "back")
r = {k: round(p[k] / reps, 4) for k in ks}
r["sum"] = round(sum(r.values()), 4)
r["span"] = round(p["span"] / reps, 4)
r["frames_per_s"] = round(S * F / (p["span"] / reps) * 1e3)
print(json.dumps(r))
