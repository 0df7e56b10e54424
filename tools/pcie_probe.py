#!/usr/bin/env python
"""Host<->device copy rates from pinned memory on this box (what bounds bench.py's e2e leg): one contiguous copy of the
per-step input, the same as 4096 rows cut into chunk-sized column blocks (cudaMemcpy2DAsync, the library's pattern), and
both directions at once."""
import json
import torch

n_streams, frames, N = 4096, 100, 480
h = torch.empty((n_streams, frames * N), dtype=torch.int16).pin_memory()
d = torch.empty_like(h, device="cuda")
ho = torch.empty((n_streams * frames * 160,), dtype=torch.uint8).pin_memory()
do = torch.empty_like(ho, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def timed(fn, reps=5):
    fn(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


out = {}
ms = timed(lambda: d.copy_(h, non_blocking=True))
out["h2d_contiguous_GBps"] = h.numel() * 2 / ms / 1e6
rt = torch.cuda.cudart()
for fc in (4, 16, 40):
    def cols():
        for f0 in range(0, frames, fc):
            w = min(fc, frames - f0) * N * 2
            rt.cudaMemcpy2DAsync(d.data_ptr() + f0 * N * 2, frames * N * 2, h.data_ptr() + f0 * N * 2, frames * N * 2, w, n_streams, 1, 0) \
                if hasattr(rt, "cudaMemcpy2DAsync") else d[:, f0 * N:(f0 + fc) * N].copy_(h[:, f0 * N:(f0 + fc) * N], non_blocking=True)
    ms = timed(cols)
    out[f"h2d_rows_x_{fc}_frames_GBps"] = h.numel() * 2 / ms / 1e6


def both():
    with torch.cuda.stream(s1): d.copy_(h, non_blocking=True)
    with torch.cuda.stream(s2): ho.copy_(do, non_blocking=True)
    torch.cuda.current_stream().wait_stream(s1); torch.cuda.current_stream().wait_stream(s2)


ms = timed(both)
out["h2d_with_d2h_ms"] = ms
out["h2d_alone_ms"] = timed(lambda: d.copy_(h, non_blocking=True))
print(json.dumps(out))
