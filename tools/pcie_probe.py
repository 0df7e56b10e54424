#!/usr/bin/env python
"""Host<->device copy rates from pinned memory on this box (what bounds bench.py's e2e leg): one contiguous copy of the
per-step input, alone and with the output's device->host copy running beside it.  (55 GB/s = 7.1 ms per step on the B200
boxes of this pool: below the 9.5 ms of kernels, so the end-to-end leg is bound by pipeline fill and drain, not by PCIe.)"""
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
def both():
    with torch.cuda.stream(s1): d.copy_(h, non_blocking=True)
    with torch.cuda.stream(s2): ho.copy_(do, non_blocking=True)
    torch.cuda.current_stream().wait_stream(s1); torch.cuda.current_stream().wait_stream(s2)


ms = timed(both)
out["h2d_with_d2h_ms"] = ms
out["h2d_alone_ms"] = timed(lambda: d.copy_(h, non_blocking=True))
print(json.dumps(out))
