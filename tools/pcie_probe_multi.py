#!/usr/bin/env python
"""Host<->device copy rates with N GPUs busy AT THE SAME TIME (one process per GPU, as bench.py runs): what bounds the
end-to-end leg when 1, 2, 4, 8 ranks copy from their own pinned buffers.  Per N: contiguous H2D, the pitched rectangles
lc3plus_enc_batch copies per chunk (4096 rows x 61 440 B out of a 96 000 B pitch, cudaMemcpy2DAsync), H2D with the
bitstream D2H beside it, and a pinned-memory memcpy on the host (what a feeder thread can do).
    python tools/pcie_probe_multi.py [max_gpus] > profiles/r2_pcie_probe.json"""
import ctypes as C
import json
import os
import subprocess
import sys
import time


def child(dev, start_at):
    import torch
    torch.cuda.set_device(dev)
    rt = None
    for name in ("libcudart.so.12", "libcudart.so"):
        try:
            rt = C.CDLL(name)
            break
        except OSError:
            continue
    if rt is None:
        import glob
        rt = C.CDLL(glob.glob(os.path.join(os.path.dirname(torch.__file__), "..", "nvidia", "cuda_runtime", "lib", "libcudart.so*"))[0])
    rt.cudaMemcpy2DAsync.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t, C.c_int, C.c_void_p]
    S, F, N = 4096, 100, 480
    h = torch.empty((S, F * N), dtype=torch.int16).pin_memory()
    h2 = torch.empty((S, F * N), dtype=torch.int16).pin_memory()
    d = torch.empty((S, F * N), dtype=torch.int16, device="cuda")
    ho = torch.empty((S * F * 160,), dtype=torch.uint8).pin_memory()
    do = torch.empty((S * F * 160,), dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    torch.cuda.synchronize()
    while time.time() < start_at:
        time.sleep(0.001)

    def timed(fn, reps=8):
        fn(); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record(); torch.cuda.synchronize()
        return a.elapsed_time(b) / reps

    out = {"dev": dev}
    nbytes = h.numel() * 2
    out["h2d_contiguous_GBps"] = nbytes / timed(lambda: d.copy_(h, non_blocking=True)) / 1e6

    def pitched():
        st = torch.cuda.current_stream().cuda_stream
        pitch = F * N * 2
        for c in range(0, F, 16):     # chunks of 16 frames
            w = min(16, F - c) * N * 2
            off = c * N * 2
            rt.cudaMemcpy2DAsync(d.data_ptr() + off, pitch, h.data_ptr() + off, pitch, w, S, 1, st)
    out["h2d_pitched_GBps"] = nbytes / timed(pitched) / 1e6

    def both():
        with torch.cuda.stream(s1):
            d.copy_(h, non_blocking=True)
        with torch.cuda.stream(s2):
            ho.copy_(do, non_blocking=True)
        torch.cuda.current_stream().wait_stream(s1); torch.cuda.current_stream().wait_stream(s2)
    ms = timed(both)
    out["h2d_plus_d2h_ms"] = ms
    out["h2d_plus_d2h_GBps"] = (nbytes + ho.numel()) / ms / 1e6
    t0 = time.perf_counter()
    for _ in range(4):
        h2.copy_(h)
    out["host_memcpy_GBps"] = 4 * nbytes / (time.perf_counter() - t0) / 1e9
    print(json.dumps(out), flush=True)


def main():
    if len(sys.argv) > 2 and sys.argv[1] == "--child":
        return child(int(sys.argv[2]), float(sys.argv[3]))
    import torch
    max_n = int(sys.argv[1]) if len(sys.argv) > 1 else torch.cuda.device_count()
    res = {"cpus": os.cpu_count(), "gpus": torch.cuda.device_count()}
    n = 1
    while n <= min(max_n, torch.cuda.device_count()):
        start = time.time() + 25.0   # children import torch and pin memory first
        ps = [subprocess.Popen([sys.executable, __file__, "--child", str(i), str(start)], stdout=subprocess.PIPE, text=True) for i in range(n)]
        rows = [json.loads(p.communicate()[0].strip().splitlines()[-1]) for p in ps]
        agg = {k: sum(r[k] for r in rows) for k in ("h2d_contiguous_GBps", "h2d_pitched_GBps", "h2d_plus_d2h_GBps", "host_memcpy_GBps")}
        res[f"n{n}"] = {"aggregate": agg, "per_gpu_min": {k: min(r[k] for r in rows) for k in agg}, "rows": rows}
        n *= 2
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
