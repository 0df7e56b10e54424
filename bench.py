#!/usr/bin/env python3
"""Headline benchmark: bit-exact LC3plus encoded frames/s on B200 (BASELINE.json metric).

Default workload (configs[1], `--config c1`): 4096 independent 48 kHz mono streams, 10 ms frames, 128 kbps, per GPU.
A "step" = one lc3plus_enc_batch_async call over FRAMES consecutive frames of every stream (c1: 100 frames = 1 s of audio
per stream, 409 600 channel-frames, 393 MB of PCM -- larger than the 126 MB L2).  `--config c3|c4|c5` time the other
BASELINE configurations the same way (c3: 12 500 x 16 kHz / 5 ms / 32 kbps per GPU = "100k streams over 8 GPUs"; c4: 8192
stereo encoders at 48 kHz / 2.5 ms / 256 kbps; c5: 3840 streams round-robin over {8, 24, 32, 44.1, 48 kHz} x {2.5, 5, 10 ms},
bitrates on an 8 kbps grid, every stream switched to a new bitrate before every step of 50 frames).

  value        channel-frames/s with PCM and bitstreams resident in HBM; the K steps are queued back to back
               (lc3plus_enc_batch_async) and waited for once; time = CUDA events on the library's streams from the first
               kernel of the first step to the end of the last, max over ranks
  e2e          same metric through the same C-ABI calls with pinned HOST buffers: H2D of the PCM and D2H of the bitstreams
               are inside the timed region (wall clock around the K calls and the final lc3plus_sync)
  verified_*   after the timed regions, the bitstreams of the LAST step of both legs are compared byte for byte with the
               unmodified reference encoder run over the same passes (oracle/_ref/refbench); a mismatch exits non-zero
  roofline     dominant kernel vs the measured HBM copy bandwidth (algorithmic bytes = PCM in + bitstream out, SURVEY.md
               section 8d), plus the integer-issue view this path is really bound by (weighted ops of the reference on this
               stimulus, profiles/wmops.json)
  cpu_baseline the unmodified ETSI reference encoder (oracle/_ref/refbench, -O2; -O3 beside it) on all host cores, bounded sample

`--impl reference` times the reference's own CPU implementation (same metric/config), rank 0 only.
Multi-GPU: one process per GPU (torchrun), each encodes its own shard (audio_codec_b200.shard.shard_range), no collective on
the data path ("scaling": "weak"); value = all ranks' frames / max-over-ranks device time.  `--single-process --gpus N` drives N
GPUs from ONE process through the C API's per-device contexts instead (no torchrun, no NCCL).
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "lc3plus_encoded_frames_per_s"
FS5 = (8000, 24000, 32000, 44100, 48000)
DMS3 = (25, 50, 100)
CONFIGS = {
    "c1": {"workload": "4096 independent 48 kHz mono streams, 10 ms frames, 128 kbps, per GPU (BASELINE configs[1])",
           "streams": 4096, "frames": 100, "wmops": "c1"},
    "c3": {"workload": "16 kHz wideband voice, 5 ms frames, 32 kbps, 12 500 mono streams per GPU (BASELINE configs[2]: 100k streams over 8 GPUs)",
           "streams": 12500, "frames": 200, "wmops": "c3"},
    "c4": {"workload": "48 kHz low-delay 2.5 ms frames, 256 kbps stereo, 8192 stereo encoders = 16 384 channel-streams per GPU (BASELINE configs[3])",
           "streams": 8192, "frames": 400, "wmops": "c4"},
    "c5": {"workload": "mixed batch: 3840 mono streams per GPU round-robin over {8, 24, 32, 44.1, 48 kHz} x {2.5, 5, 10 ms}, bitrates on an 8 kbps "
                       "grid, lc3_enc_set_bitrate on every stream before every step of 50 frames (BASELINE configs[4])",
           "streams": 3840, "frames": 50, "wmops": None},
}


def frame_len(fs, dms):
    return (48000 if fs == 44100 else fs) // 100 * dms // 100


def legal_bitrates(fs, dms, ch=1):
    """8 kbps grid inside the legal range of the frame size (ETSI/Readme.txt:277-284; fx/setup_enc_lc3.c:137-149)."""
    k = (441.0 / 480) if fs == 44100 else 1.0
    lo = int(20 * 8.0 * (10000.0 / dms) * k) * ch
    hi = int(400 * 8.0 * (10000.0 / dms) * k) * ch
    return np.arange((lo + 7999) // 8000 * 8000, hi + 1, 8000)


class Workload:
    """Per-stream configuration, PCM rows and byte strides of one rank's shard of a BASELINE configuration."""

    def __init__(self, name, first_stream, n_streams, frames):
        from audio_codec_b200.synth import make_streams, make_streams_fast
        self.name, self.frames, self.S = name, frames, n_streams
        F = frames
        if name == "c5":
            self.spec = []
            for i in range(n_streams):
                s = first_stream + i
                fs, dms = FS5[s % 5], DMS3[(s // 5) % 3]
                rng = np.random.RandomState(99 + s)
                self.spec.append((fs, dms, 1, 16, [int(b) for b in rng.choice(legal_bitrates(fs, dms), 8)]))
            self.stride = F * 480 * 2
            self.pcm = np.zeros((n_streams, self.stride // 2), dtype=np.int16)
            base = {}
            for i, (fs, dms, ch, bits, brs) in enumerate(self.spec):    # 16 distinct signals per (fs, dms), rolled per stream
                N = frame_len(fs, dms)
                key = (fs, dms)
                if key not in base:
                    base[key] = make_streams(16, F * N, fs, N, first_stream=5 * DMS3.index(dms) + FS5.index(fs))
                self.pcm[i, :F * N] = np.roll(base[key][(i // 15) % 16], 97 * (i // 240))
        else:
            fs, ch, dms, br = {"c1": (48000, 1, 100, 128000), "c3": (16000, 1, 50, 32000), "c4": (48000, 2, 25, 256000)}[name]
            N = frame_len(fs, dms)
            self.spec = [(fs, dms, ch, 16, [br])] * n_streams
            mono = make_streams_fast(n_streams * ch, F * N, fs, N)
            if first_stream:
                mono = np.roll(mono, 131 * (first_stream // max(n_streams, 1)), axis=1)
            self.pcm = mono.reshape(n_streams, ch * F * N)     # encoder s: channels planar = mono rows s*ch .. s*ch + ch - 1
            self.stride = ch * F * N * 2
        self.max_nb = max(max(br) * frame_len(fs, dms) // (8 * fs) for fs, dms, ch, bits, br in self.spec)
        self.ostride = F * self.max_nb
        self.channel_frames = sum(ch for fs, dms, ch, bits, br in self.spec) * F
        self.passes_per_bitrate = 1

    def num_bytes(self, i, p):
        fs, dms, ch, bits, br = self.spec[i]
        return br[p % len(br)] * frame_len(fs, dms) // (8 * fs)

    def algorithmic_bytes(self, p=0):
        """PCM bytes in + bitstream bytes out of one step"""
        F = self.frames
        return sum(ch * F * frame_len(fs, dms) * 2 + F * (br[p % len(br)] * frame_len(fs, dms) // (8 * fs)) for fs, dms, ch, bits, br in self.spec)

    def bitrates(self, p):
        return np.array([br[p % len(br)] for fs, dms, ch, bits, br in self.spec], dtype=np.int32)

    def switching(self):
        return any(len(br) > 1 for fs, dms, ch, bits, br in self.spec)

    def make_encoders(self, ac):
        encs = []
        for fs, dms, ch, bits, br in self.spec:
            encs.append(ac.Encoder(fs, ch, dms, br[0]))
        return ac.EncoderArray(encs)


def refbench_path(o3=False):
    p = os.path.join(ROOT, "oracle", "_ref", "refbench_O3" if o3 else "refbench")
    return p if os.path.exists(p) else None


def run_refbench(w, idx, threads, repeat, want_out=False, o3=False):
    """Unmodified reference encoder over streams idx of workload w, `repeat` passes (bitrate of pass p = schedule[p]).
    Returns (channel-frames/s, seconds, out rows or None)."""
    with tempfile.TemporaryDirectory(dir="/tmp") as d:
        np.ascontiguousarray(w.pcm[idx]).tofile(os.path.join(d, "pcm.raw"))
        with open(os.path.join(d, "spec.txt"), "w") as f:
            f.write(f"{len(idx)} {w.stride} {w.ostride}\n")
            for i in idx:
                fs, dms, ch, bits, br = w.spec[i]
                f.write(f"{fs} {dms} {ch} {bits} {len(br)} {' '.join(map(str, br))}\n")
        outp = os.path.join(d, "out.bin") if want_out else "-"
        o = subprocess.run([refbench_path(o3), "@" + os.path.join(d, "spec.txt"), os.path.join(d, "pcm.raw"), str(w.frames),
                            str(max(1, min(threads, len(idx)))), outp, str(repeat)], capture_output=True, text=True, check=True).stdout.split()
        out = np.fromfile(outp, dtype=np.uint8).reshape(len(idx), w.ostride) if want_out else None
    return float(o[2]), float(o[1]), out


def verify_against_reference(w, out_rows, passes, what):
    """Byte-compare some streams of `out_rows` [S, ostride] (the bitstreams of pass number `passes` over the PCM, state carried
    from pass to pass) with the unmodified reference encoder run over the same passes.  Returns (streams, frames) verified; a
    mismatch ends the benchmark with a non-zero exit code."""
    if refbench_path() is None:
        return 0, 0
    cand = (0, 7, 14, 15, 31, 32, 63, 64, 1000, 2047, w.S - 2, w.S - 1) if w.name != "c5" else tuple(range(15)) + (239, 240, 1000, w.S - 1)
    idx = sorted({s for s in cand if 0 <= s < w.S})
    _, _, want = run_refbench(w, idx, os.cpu_count() or 1, passes, want_out=True)
    bad = []
    for k, i in enumerate(idx):
        n = w.frames * w.num_bytes(i, passes - 1)
        if not np.array_equal(out_rows[i, :n], want[k, :n]):
            bad.append(i)
    if bad:
        print(f"bench.py: {what}: output differs from the reference encoder for streams {bad[:8]}", file=sys.stderr)
        sys.exit(3)
    return len(idx), len(idx) * w.frames


def clock_sampler(stop, rows, dev):
    """SM clock and throttle reasons during the timed region, sampled in-process through NVML (spawning
    nvidia-smi next to the timed loop stalls kernel launches for hundreds of ms)."""
    try:
        import pynvml as nv
        nv.nvmlInit()
        h = nv.nvmlDeviceGetHandleByIndex(dev)
        mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
        R = {"hw_slowdown": nv.nvmlClocksThrottleReasonHwSlowdown, "hw_thermal_slowdown": nv.nvmlClocksThrottleReasonHwThermalSlowdown,
             "sw_thermal_slowdown": nv.nvmlClocksThrottleReasonSwThermalSlowdown, "sw_power_cap": nv.nvmlClocksThrottleReasonSwPowerCap}
        while not stop.is_set():
            bits = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
            rows.append([str(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)), str(mx)] +
                        ["Active" if bits & R[k] else "Not Active" for k in ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")])
            stop.wait(0.05)
        return
    except Exception:
        pass
    q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    while not stop.is_set():
        try:
            o = subprocess.run(["nvidia-smi", "-i", str(dev), f"--query-gpu={q}", "--format=csv,noheader,nounits"],
                               capture_output=True, text=True, timeout=5).stdout.strip().split(",")
            rows.append([x.strip() for x in o])
        except Exception:
            pass
        stop.wait(0.5)


def summarize_clocks(rows):
    sm = [int(r[0]) for r in rows if r and r[0].isdigit()]
    mx = [int(r[1]) for r in rows if len(r) > 1 and r[1].isdigit()]
    names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
    reasons = [n for i, n in enumerate(names) if any(len(r) > 2 + i and r[2 + i].startswith("Active") for r in rows)]
    return {"sm_mhz": int(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
            "samples": len(sm)}


def sample_indices(w, cores):
    """streams of the bounded CPU sample: the first ones of the shard (c5: a whole number of 15-stream rounds, so every
    (sample rate, frame size) pair is in it)"""
    ns = int(min(w.S, max(4 * cores, 64)))
    if w.name == "c5":
        ns = min(w.S, max(15, ns // 15 * 15))
    return list(range(ns))


def cpu_reference_rate(w, target_s, o3=False):
    """Bounded sample of the workload on all host cores: calibrate, then one run of about target_s seconds."""
    cores = os.cpu_count() or 1
    idx = sample_indices(w, cores)
    nbr = max(len(s[4]) for s in w.spec)
    rate, secs, _ = run_refbench(w, idx, cores, nbr, o3=o3)
    rep = int(max(nbr, min(400, round(target_s / max(secs / nbr, 1e-3)))))
    if rep > nbr:
        rate, secs, _ = run_refbench(w, idx, cores, rep, o3=o3)
    cf = sum(w.spec[i][2] for i in idx) * w.frames * rep
    sample = f"{len(idx)} of the workload's streams x {w.frames * rep} frames ({cf} channel-frames, {secs:.1f} s) on {cores} threads, gcc {'-O3' if o3 else '-O2'}"
    return rate, cores, sample


def config_dict(name, w, world, S):
    d = {"workload": CONFIGS[name]["workload"], "config": name, "streams_per_gpu": S, "frames_per_step": w.frames,
         "channel_frames_per_step": w.channel_frames * world,
         "l2": "per-step input %d MB > 126 MB L2 (no flush needed)" % (w.pcm.nbytes // 1000000)}
    return d


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cfg = CONFIGS[args.config]
    S, F = args.streams or cfg["streams"], args.frames or cfg["frames"]
    cores = os.cpu_count() or 1
    # the sample's streams are the first ones of the workload: build only those
    ns = int(min(S, max(4 * cores, 64)))
    if args.config == "c5":
        ns = min(S, max(15, ns // 15 * 15))
    w = Workload(args.config, 0, ns, F)
    full = {"S": S}
    line = {"impl": "reference", "metric": METRIC, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int16/int32 fixed point",
            "data": "synthetic"}
    wd = config_dict(args.config, w, args.gpus, S)
    wd["channel_frames_per_step"] = int(w.channel_frames / ns * S) * args.gpus
    wd["l2"] = "per-step input %d MB > 126 MB L2 (no flush needed)" % (int(w.pcm.nbytes / ns * S) // 1000000)
    line["config"] = wd
    if refbench_path() is None:
        line["unavailable"] = "oracle/_ref/refbench not built (reference tree absent at build time)"
        print(json.dumps(line))
        return
    idx = list(range(ns))
    nbr = max(len(s[4]) for s in w.spec)
    rate, secs, _ = run_refbench(w, idx, cores, nbr)                      # calibration (also a warm-up)
    per_pass = max(secs / nbr, 1e-3)
    total_steps = args.steps + args.warmup
    step_s = min(3.0, 240.0 / total_steps)                                 # ~3 s per step, the whole run within a few minutes
    rep = int(max(nbr, min(400, round(step_s / per_pass))))
    for _ in range(args.warmup):
        run_refbench(w, idx, cores, rep)
    frames = 0.0
    t = 0.0
    for _ in range(args.steps):
        r, s, _ = run_refbench(w, idx, cores, rep)
        frames += r * s
        t += s
    value = frames / t
    sample = (f"each step = {ns} of the workload's streams x {F * rep} frames on {cores} host threads (unmodified ETSI reference, gcc -O2)")
    line.update({"value": value, "ms_per_step": 1e3 * t / args.steps,
                 "cpu_baseline": {"value": value, "unit": "frames/s", "cores": cores, "kind": "reference", "sample": sample},
                 "e2e": {"value": value, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                 "realtime_channels": value / (1000.0 / (w.spec[0][1] / 10.0)) if args.config != "c5" else None,
                 "sample_channel_frames_per_step": int(w.channel_frames * rep)})
    del full
    print(json.dumps(line))


KS = ("stage", "fir", "scalar", "pitch", "mdct", "sns", "tns", "rate", "quant", "back")


class Unit:
    """one GPU's shard: workload, encoders (living on that device), device and pinned host buffers, pass counter"""

    def __init__(self, ac, torch, name, dev, shard, n_shards, S, F, total=0):
        from audio_codec_b200.shard import shard_range
        lo, hi = shard_range(total if total else S * n_shards, shard, n_shards)
        self.dev = dev
        self.w = Workload(name, lo, hi - lo, F)
        torch.cuda.set_device(dev)
        ac.set_device(dev)
        self.encs = self.w.make_encoders(ac)
        ac.set_device(-1)
        self.d_pcm = torch.from_numpy(self.w.pcm).to(f"cuda:{dev}")
        self.d_out = torch.zeros((self.w.S, self.w.ostride), dtype=torch.uint8, device=f"cuda:{dev}")
        self.h_pcm = self.h_out = None
        self.p = 0          # passes over the PCM so far (state carries on; c5: selects the bitrate schedule entry)

    def queue(self, ac, host=False):
        w = self.w
        if w.switching():
            ac.set_bitrate_many(self.encs, w.bitrates(self.p))   # part of the workload: mid-stream switching before every step
        src, dst = (self.h_pcm, self.h_out) if host else (self.d_pcm, self.d_out)
        ac.enc_batch(self.encs, w.frames, src.data_ptr(), w.stride, 16, dst.data_ptr(), w.ostride, wait=False)
        self.p += 1


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--config", default="c1", choices=sorted(CONFIGS))
    ap.add_argument("--streams", type=int, default=0, help="streams per GPU (0: the configuration's)")
    ap.add_argument("--frames", type=int, default=0, help="frames per stream per step (0: the configuration's)")
    ap.add_argument("--chunk", type=int, default=0, help="channel-frames per internal chunk (0: library default); profiling aid")
    ap.add_argument("--total-streams", type=int, default=0,
                    help="strong scaling: this many streams in TOTAL, sharded over the GPUs with shard_range (e.g. --config c3 --total-streams 100000)")
    ap.add_argument("--single-process", action="store_true", help="drive --gpus N devices from this one process (no torchrun)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-decoder", action="store_true", help="skip the short decoder leg (key \"decoder\" of the c1 line)")
    ap.add_argument("--no-alone", action="store_true", help="skip the kernels-alone pass (roofline.kernel_ms_alone)")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        return reference_arm(args)

    import torch
    import audio_codec_b200 as ac
    cfg = CONFIGS[args.config]
    S, F = args.streams or cfg["streams"], args.frames or cfg["frames"]
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU path)"
    dist = None
    if args.single_process:
        rank, world, n_shards = 0, 1, args.gpus
        assert torch.cuda.device_count() >= args.gpus, "--single-process --gpus N needs N visible devices"
        units = [Unit(ac, torch, args.config, d, d, n_shards, S, F, args.total_streams) for d in range(args.gpus)]
        torch.cuda.set_device(0)
    else:
        n_shards = world
        torch.cuda.set_device(local)
        if world > 1:
            import torch.distributed as dist
            dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        units = [Unit(ac, torch, args.config, local, rank, world, S, F, args.total_streams)]
    u0 = units[0]

    def barrier():
        if dist is not None:
            dist.barrier()
        for u in units:
            torch.cuda.synchronize(u.dev)

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    if args.chunk:
        ac.set_chunk_frames(args.chunk)

    def steps(k, host=False):
        for _ in range(k):
            for u in units:
                u.queue(ac, host)
        ac.sync()

    steps(args.warmup)
    rows, stop = [], threading.Event()
    th = threading.Thread(target=clock_sampler, args=(stop, rows, u0.dev), daemon=True)
    if rank == 0:
        th.start()
    barrier()
    l0 = ac.kernel_launches()
    ac.profile_enable(1)                       # device span only: the calls stay asynchronous
    t0 = time.perf_counter()
    steps(args.steps)
    for u in units:
        torch.cuda.synchronize(u.dev)
    wall = time.perf_counter() - t0
    spans = [ac.profile_read_device(u.dev)["span"] for u in units]
    ac.profile_enable(0)
    launches = ac.kernel_launches() - l0
    last_out = u0.d_out.cpu().numpy() if rank == 0 else None     # bitstreams of the last timed step, checked after the timing legs
    passes_dev = u0.p
    barrier()
    dev_s = max_over_ranks(max(spans) / 1e3)
    wall_s = max_over_ranks(wall)
    cf_step = sum(u.w.channel_frames for u in units)             # this process's channel-frames per step
    if dist is not None and args.total_streams:                  # shards may differ by one stream: sum over the ranks
        t = torch.tensor([float(cf_step)], dtype=torch.float64, device="cuda")
        dist.all_reduce(t)
        cf_step = t.item() / world
    frames_total = float(cf_step) * args.steps * world
    value = frames_total / dev_s

    # per-kernel times with every kernel ALONE on the device (the chunk's kernels strictly one after the other, one chunk per
    # step and bucket): in the timed region kernels of neighbouring chunks share the SMs
    alone = None
    if not args.no_alone:
        os.environ["LC3PLUS_B200_CHAIN"] = "1"
        ac.set_chunk_frames(1 << 26)
        u0.queue(ac); ac.sync()
        ac.profile_enable(2)
        for _ in range(3):
            u0.queue(ac); ac.sync()
        alone = ac.profile_read_device(u0.dev)
        ac.profile_enable(0)
        del os.environ["LC3PLUS_B200_CHAIN"]
        ac.set_chunk_frames(args.chunk if args.chunk else 0)   # 0: the library's defaults
    barrier()

    # end to end: pinned host PCM in, pinned host bitstreams out, same calls
    e2e = None
    if not args.no_e2e:
        for u in units:
            torch.cuda.set_device(u.dev)
            u.h_pcm = torch.from_numpy(u.w.pcm).pin_memory()
            u.h_out = torch.zeros((u.w.S, u.w.ostride), dtype=torch.uint8).pin_memory()
        torch.cuda.set_device(u0.dev)
        ksteps = max(3, min(args.steps, 10))
        steps(2, host=True)   # warm-up (allocates both staging buffers)
        barrier()
        t0 = time.perf_counter()
        steps(ksteps, host=True)
        e2e_s = max_over_ranks(time.perf_counter() - t0)
        barrier()
        e2e = {"value": float(cf_step) * ksteps * world / e2e_s, "unit": "frames/s",
               "h2d_bytes_per_step": int(sum(u.w.pcm.nbytes for u in units)) * world,
               "d2h_bytes_per_step": int(sum(u.w.algorithmic_bytes() - u.w.pcm.nbytes for u in units)) * world if not u0.w.switching()
               else int(sum(u.w.S * u.w.ostride for u in units)) * world, "steps": ksteps,
               "note": "pinned host buffers -> lc3plus_enc_batch_async x steps, lc3plus_sync -> pinned host buffers; copies inside the timed region (wall clock)"}
    stop.set()
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    # the timed bytes are the reference encoder's bytes: device-resident leg (last timed step) and end-to-end leg (last step)
    v_streams, v_frames = verify_against_reference(u0.w, last_out, passes_dev, "device-resident leg")
    verified = {"verified_streams": v_streams, "verified_frames": v_frames,
                "verified_against": "unmodified ETSI reference encoder (oracle/_ref/refbench) over the same %d passes of the PCM, a sample of "
                                    "rank 0's streams (first, last, tile edges, silence-leading ones)" % passes_dev}
    if e2e is not None:
        e2e["verified_streams"], e2e["verified_frames"] = verify_against_reference(u0.w, u0.h_out.numpy(), u0.p, "end-to-end leg")

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak, peak_src = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json)") if "hbm_gbs" in peaks else (6650.0, "fallback (B200_PROFILING.md)")
    sm_clock = summarize_clocks(rows)
    roofline = None
    if alone is not None:
        kern = {k: alone[k] / alone["calls"] for k in KS if alone[k] > 0}   # ms per step of this rank's shard, kernel alone
        top = max(kern, key=kern.get)
        avg_ms = kern[top]
        algo = u0.w.algorithmic_bytes(u0.p - 1)
        achieved = algo / (avg_ms * 1e-3) / 1e9
        traffic = None
        try:
            # DRAM bytes of the dominant kernel from the committed `ncu --set full` capture of this configuration
            # (tools/ncu_summary.py), per launch, rescaled linearly to this run's channel-frames per launch
            tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
            if tj.get("_config", "c1") == args.config:
                traffic = tj[f"k_{top}"][0]["dram_bytes"] * u0.w.channel_frames / tj["_frames_per_launch"]
        except Exception:
            pass
        int_peak, int_src = 148 * 4 * 32 * (sm_clock["sm_mhz"] or 1965) * 1e6, "148 SMs x 4 schedulers x 32 lanes x median SM clock"
        try:   # measured on this pool's B200 by tools/int_peak.cu (mixed IMAD+LOP3 stream, 97 % of the issue limit)
            ip = json.load(open(os.path.join(ROOT, "profiles", "int_peak.json")))
            int_peak = ip["mixed_thread_instr_per_s"] * (sm_clock["sm_mhz"] or 1965) * 1e3 / ip["sm_clock_khz_max"]
            int_src = "measured: profiles/int_peak.json (tools/int_peak.cu), scaled to the median SM clock of this run"
        except Exception:
            pass
        wops = None
        try:
            wops = json.load(open(os.path.join(ROOT, "profiles", "wmops.json")))[cfg["wmops"]]["weighted_ops_per_channel_frame"]
        except Exception:
            pass
        per_gpu = value / (world * len(units))
        roofline = {"bound": "hbm", "kernel": f"k_{top}", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                    "frac": achieved / hbm_peak, "traffic": traffic, "peak_source": peak_src,
                    "algorithmic_bytes_per_launch": algo, "algorithmic_bytes_per_channel_frame": algo / u0.w.channel_frames, "avg_launch_ms": avg_ms,
                    "timing": "CUDA events on the launching stream, each kernel alone on the device (3 steps after the timed region, one "
                              "launch of each kernel per step and (sample rate, frame size) bucket; %d channel-frames per step); peak is the burst figure" % u0.w.channel_frames,
                    "kernel_ms_alone": {k: round(v, 4) for k, v in kern.items()},
                    "kernel_ms_share": {k: v / sum(kern.values()) for k, v in kern.items()},
                    "int_alu": {"note": "the path is integer-issue bound, not HBM bound (SURVEY.md 8d): reference weighted ops per channel-frame on this "
                                        "stimulus (profiles/wmops_%s.txt) x channel-frames/s per GPU over the integer issue peak" % cfg["wmops"],
                                "peak_source": int_src, "weighted_ops_per_channel_frame": wops,
                                "achieved_wops_per_s": per_gpu * wops if wops else None,
                                "peak_thread_instr_per_s": int_peak, "frac": per_gpu * wops / int_peak if wops else None}}
    audio_s = sum(ch * u.w.frames * dms / 10000.0 for u in units for fs, dms, ch, bits, br in u.w.spec)   # seconds of audio per step
    line = {"metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world * len(units), "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * dev_s / args.steps, "higher_is_better": True, "scaling": "strong" if args.total_streams else "weak", "vs_baseline": None,
            "dtype": "int16/int32 fixed point", "data": "synthetic",
            "config": config_dict(args.config, u0.w, world * len(units), S),
            "realtime_channels": audio_s * world * args.steps / dev_s, "wall_ms_per_step": 1e3 * wall_s / args.steps,
            "e2e": e2e, "gpu_launches": int(launches), "clocks": sm_clock, "roofline": roofline}
    if args.single_process:
        line["processes"] = 1
    if args.total_streams:
        line["config"]["total_streams"] = args.total_streams
        line["config"]["streams_per_gpu"] = u0.w.S
    line.update(verified)
    if world == 1 and len(units) == 1 and not args.no_cpu_baseline and refbench_path() is not None:
        rate, cores, sample = cpu_reference_rate(u0.w, 12.0)
        line["cpu_baseline"] = {"value": rate, "unit": "frames/s", "cores": cores, "kind": "reference", "sample": sample}
        if refbench_path(o3=True) is not None:
            r3, _, s3 = cpu_reference_rate(u0.w, 6.0, o3=True)
            line["cpu_baseline"]["value_O3"] = r3
            line["cpu_baseline"]["sample_O3"] = s3
    if world == 1 and len(units) == 1 and args.config == "c1" and not args.no_decoder:
        try:   # a short leg of the decoder (row f4) beside the headline, which stays the encoder's; never fatal for the line
            import bench_dec
            line["decoder"] = bench_dec.quick_leg()
        except Exception as e:   # noqa: BLE001
            line["decoder"] = {"error": repr(e)[:200]}
    print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
