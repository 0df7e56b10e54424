#!/usr/bin/env python
"""Per-warp (or per-block) run times of the kernels, from a library built with -DLC3_DBG_TIMING
(make -C audio_codec_b200/csrc clean && make -C audio_codec_b200/csrc NVFLAGS+=-DLC3_DBG_TIMING; see lc3_dbg.cuh).

Runs BASELINE configs[1] (4096 streams x 100 frames, one chunk), then prints for every instrumented kernel the
distribution of the work units' run times, the launch span, how busy the SM slots would be if every unit took the
median, and the units that ran longer than three medians -- the stragglers that ncu's averages hide.
Usage (on the GPU box):  python tools/warp_times.py [--bursts] [out.npz]   (--bursts: transient-rich stimulus)
"""
import ctypes, sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import audio_codec_b200 as ac
from audio_codec_b200 import api
from audio_codec_b200.synth import make_streams_fast, make_bursts

ROWS, COLS, KERNELS = 16384, 8, 4
UNITS = {"front": ["k_tns (warp = tile-frame)", "k_mdct (block = tile-frame)"], "sns": ["k_sns (warp = tile-frame)"],
         "back": ["k_quant (block = tile-frame)", "k_back (warp = tile-frame)"], "stream": ["k_pitch (warp = stream)", "k_rate (warp = stream)"]}

def main():
    S, F, N = 4096, 100, 480
    bursts = '--bursts' in sys.argv
    if bursts: sys.argv.remove('--bursts')
    pcm = make_bursts(S, F * N, 48000) if bursts else make_streams_fast(S, F * N, 48000, N)
    encs = [ac.Encoder(48000, 1, 100, 128000) for _ in range(S)]
    ac.set_chunk_frames(1 << 30)
    d_pcm = torch.from_numpy(pcm).cuda()
    d_out = torch.empty((S, F * 160), dtype=torch.uint8, device="cuda")
    for _ in range(2):
        ac.enc_batch(encs, F, d_pcm.data_ptr(), F * N * 2, 16, d_out.data_ptr(), F * 160)
    torch.cuda.synchronize()
    lib = api.load_library()
    out = {}
    for unit, names in UNITS.items():
        fn = getattr(lib, "lc3k_dbg_read_" + unit, None)
        if fn is None:
            sys.exit("library was not built with -DLC3_DBG_TIMING")
        buf = np.zeros(KERNELS * ROWS * COLS, dtype=np.uint64)
        rc = fn(buf.ctypes.data_as(ctypes.c_void_p))
        assert rc == 0, rc
        buf = buf.reshape(KERNELS, ROWS, COLS).astype(np.int64)
        for k, name in enumerate(names):
            a = buf[k]
            a = a[(a[:, 0] > 0) & (a[:, 5] > 0)]
            out[name.split()[0]] = a
            if not len(a):
                print(f"{name}: no rows"); continue
            t0, t1 = a[:, 0], a[:, 5]
            dur = (t1 - t0) / 1e3
            span = (t1.max() - t0.min()) / 1e3
            med = np.median(dur)
            slow = dur > 3 * med
            print(f"{name}: {len(a)} units, span {span:.0f} us; unit run time us: min {dur.min():.0f} p10 {np.percentile(dur,10):.0f} "
                  f"median {med:.0f} mean {dur.mean():.0f} p90 {np.percentile(dur,90):.0f} p99 {np.percentile(dur,99):.0f} max {dur.max():.0f}; "
                  f"{slow.sum()} units over 3x median, their time {dur[slow].sum() / dur.sum() * 100:.1f}% of the total")
    if len(sys.argv) > 1:
        np.savez_compressed(sys.argv[1], **out)

if __name__ == "__main__":
    main()
