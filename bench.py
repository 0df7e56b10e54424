#!/usr/bin/env python3
"""Headline benchmark: bit-exact LC3plus encoded frames/s on B200 (BASELINE.json metric).

Workload (configs[1]): 4096 independent 48 kHz mono streams, 10 ms frames, 128 kbps, per GPU.
A "step" = one lc3plus_enc_batch call over FRAMES consecutive frames of every stream (default 100 frames =
1 s of audio per stream, 409 600 channel-frames, 393 MB of PCM -- larger than the 126 MB L2).

  value        frames/s with PCM and bitstreams resident in HBM, timed with CUDA events on the library's
               launching stream (first kernel start -> last kernel end of each call), max over ranks
  e2e          same metric through the same C-ABI call with pinned HOST buffers: H2D of the PCM and D2H of
               the bitstreams are inside the timed region
  roofline     dominant kernel vs the measured HBM copy bandwidth (algorithmic bytes = PCM in + bitstream out,
               SURVEY.md section 8d), plus the integer-ALU view this path is really bound by
  cpu_baseline the unmodified ETSI reference encoder (oracle/_ref/refbench) on all host cores, bounded sample

`--impl reference` times the reference's own CPU implementation (same metric/config), rank 0 only.
Multi-GPU: one process per GPU (torchrun), each encodes its own 4096-stream shard, no collective on the data
path ("scaling": "weak"); value = all ranks' frames / max-over-ranks device time.
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

FS, DMS, BITRATE = 48000, 100, 128000
N = 480
NBYTES = BITRATE * N // (8 * FS)
ALGO_BYTES_PER_FRAME = N * 2 + NBYTES          # 960 B PCM in + 160 B bitstream out
WOPS_PER_FRAME = 1.22e5                        # reference's own WMOPS counter at this config
METRIC = "lc3plus_encoded_frames_per_s"
WORKLOAD = "4096 independent 48 kHz mono streams, 10 ms frames, 128 kbps, per GPU (BASELINE configs[1])"


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


def refbench_path():
    p = os.path.join(ROOT, "oracle", "_ref", "refbench")
    return p if os.path.exists(p) else None


def run_refbench(pcm_block, n_streams, n_frames, threads, repeat):
    """Unmodified reference encoder over pcm_block [n_streams, n_frames*N] int16; returns (frames/s, seconds)."""
    with tempfile.NamedTemporaryFile(suffix=".raw", dir="/tmp") as f:
        pcm_block.tofile(f.name)
        o = subprocess.run([refbench_path(), f.name, str(FS), str(DMS), str(BITRATE), str(n_streams), str(n_frames),
                            str(threads), "-", str(repeat)], capture_output=True, text=True, check=True).stdout.split()
    return float(o[2]), float(o[1])


VERIFY_STREAMS = (0, 7, 31, 32, 63, 64, 1000, 2047, 4094, 4095)   # 7 and 63 open with 50 frames of digital silence


def verify_against_reference(pcm, out, frames, repeat, what):
    """Byte-compare some streams of `out` [S, frames, NBYTES] (the bitstreams of the `repeat`-th pass over `pcm`, state carried
    from pass to pass) with the unmodified reference encoder run over the same passes (oracle/_ref/refbench ... out.bin).
    Returns (streams, frames) verified; a mismatch ends the benchmark with a non-zero exit code."""
    if refbench_path() is None:
        return 0, 0
    idx = sorted({s for s in VERIFY_STREAMS if s < pcm.shape[0]})
    block = np.ascontiguousarray(pcm[idx, :frames * N])
    with tempfile.TemporaryDirectory(dir="/tmp") as d:
        block.tofile(os.path.join(d, "pcm.raw"))
        subprocess.run([refbench_path(), os.path.join(d, "pcm.raw"), str(FS), str(DMS), str(BITRATE), str(len(idx)), str(frames),
                        str(min(len(idx), os.cpu_count() or 1)), os.path.join(d, "out.bin"), str(repeat)], capture_output=True, check=True)
        want = np.fromfile(os.path.join(d, "out.bin"), dtype=np.uint8).reshape(len(idx), frames, NBYTES)
    got = np.asarray(out)[idx]
    if not np.array_equal(got, want):
        bad = [(idx[i], int(np.argmax((got[i] != want[i]).any(axis=1)))) for i in range(len(idx)) if not np.array_equal(got[i], want[i])]
        print(f"bench.py: {what}: output differs from the reference encoder at (stream, first frame) {bad[:8]}", file=sys.stderr)
        sys.exit(3)
    return len(idx), len(idx) * frames


def cpu_reference_rate(pcm, frames, target_s):
    """Bounded sample of the workload on all host cores: calibrate, then one run of about target_s seconds."""
    cores = os.cpu_count() or 1
    ns = int(min(pcm.shape[0], max(4 * cores, 64)))
    block = np.ascontiguousarray(pcm[:ns, :frames * N])
    rate, secs = run_refbench(block, ns, frames, cores, 1)
    rep = int(max(1, min(200, round(target_s / max(secs, 1e-3)))))
    if rep > 1:
        rate, secs = run_refbench(block, ns, frames, cores, rep)
    sample = f"{ns} of the workload's streams x {frames * rep} frames ({ns * frames * rep} frames, {secs:.1f} s) on {cores} threads"
    return rate, cores, sample


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from audio_codec_b200.synth import make_streams_fast
    line = {"impl": "reference", "metric": METRIC, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int16/int32 fixed point",
            "data": "synthetic", "config": {"workload": WORKLOAD, "frames_per_step_sample": None}}
    if refbench_path() is None:
        line["unavailable"] = "oracle/_ref/refbench not built (reference tree absent at build time)"
        print(json.dumps(line))
        return
    cores = os.cpu_count() or 1
    ns = int(min(args.streams, max(4 * cores, 64)))
    pcm = make_streams_fast(ns, args.frames * N, FS, N)
    rate, secs = run_refbench(pcm, ns, args.frames, cores, 1)            # calibration (also a warm-up)
    rep = int(max(1, min(100, round(3.0 / max(secs, 1e-3)))))             # ~3 s per step
    total_steps = args.steps + args.warmup
    budget = 240.0
    if total_steps * 3.0 > budget:
        rep = int(max(1, rep * budget / (total_steps * 3.0)))
    for _ in range(args.warmup):
        run_refbench(pcm, ns, args.frames, cores, rep)
    frames = 0.0
    t = 0.0
    for _ in range(args.steps):
        r, s = run_refbench(pcm, ns, args.frames, cores, rep)
        frames += r * s
        t += s
    value = frames / t
    sample = f"each step = {ns} of the workload's streams x {args.frames * rep} frames on {cores} host threads (unmodified ETSI reference, gcc -O2)"
    line.update({"value": value, "ms_per_step": 1e3 * t / args.steps,
                 "cpu_baseline": {"value": value, "unit": "frames/s", "cores": cores, "kind": "reference", "sample": sample},
                 "e2e": {"value": value, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                 "realtime_channels": value / 100.0})
    line["config"]["frames_per_step_sample"] = ns * args.frames * rep
    print(json.dumps(line))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--streams", type=int, default=4096, help="streams per GPU")
    ap.add_argument("--frames", type=int, default=100, help="frames per stream per step")
    ap.add_argument("--chunk", type=int, default=0, help="channel-frames per internal chunk (0: library default); profiling aid")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        return reference_arm(args)

    import torch
    import audio_codec_b200 as ac
    from audio_codec_b200.synth import make_streams_fast
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (there is no CPU path)"
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    S, F = args.streams, args.frames
    pcm = make_streams_fast(S, F * N, FS, N)                 # every rank: its own shard of S streams
    if rank:
        pcm = np.roll(pcm, 131 * rank, axis=1)
    encs = ac.EncoderArray([ac.Encoder(FS, 1, DMS, BITRATE) for _ in range(S)])  # the C call's `encoders` array, built once
    if args.chunk:
        ac.set_chunk_frames(args.chunk)
    d_pcm = torch.from_numpy(pcm).cuda()
    d_out = torch.zeros((S, F, NBYTES), dtype=torch.uint8, device="cuda")

    def step_dev():
        ac.enc_batch(encs, F, d_pcm.data_ptr(), F * N * 2, 16, d_out.data_ptr(), F * NBYTES)

    for _ in range(args.warmup):
        step_dev()
    rows, stop = [], threading.Event()
    th = threading.Thread(target=clock_sampler, args=(stop, rows, local), daemon=True)
    if rank == 0:
        th.start()
    barrier()
    l0 = ac.kernel_launches()
    ac.profile_enable(True)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_dev()
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    prof = ac.profile_read()
    ac.profile_enable(False)
    launches = ac.kernel_launches() - l0
    last_out = d_out.cpu().numpy() if rank == 0 else None     # bitstreams of the last timed step, checked after the timing legs
    passes = args.warmup + args.steps                          # every step encodes the same PCM again, state carried
    barrier()
    dev_s = max_over_ranks(prof["span"] / 1e3)
    wall_s = max_over_ranks(wall)
    frames_total = float(S) * F * args.steps * world
    value = frames_total / dev_s

    # per-kernel times with every kernel ALONE on the device (the chunk's kernels strictly one after the other, one chunk per
    # step): in the timed region above kernels of neighbouring chunks share the SMs, so their event brackets overlap
    os.environ["LC3PLUS_B200_CHAIN"] = "1"
    ac.set_chunk_frames(S * F)
    step_dev()
    torch.cuda.synchronize()
    ac.profile_enable(True)
    for _ in range(3):
        step_dev()
    torch.cuda.synchronize()
    passes += 4
    alone = ac.profile_read()
    ac.profile_enable(False)
    del os.environ["LC3PLUS_B200_CHAIN"]
    ac.set_chunk_frames(args.chunk if args.chunk else 0)   # 0: the library's defaults
    barrier()

    # end to end: pinned host PCM in, pinned host bitstreams out, same call
    e2e = None
    if not args.no_e2e:
        h_pcm = torch.from_numpy(pcm).pin_memory()
        h_out = torch.zeros((S, F, NBYTES), dtype=torch.uint8).pin_memory()
        ksteps = max(3, min(args.steps, 10))
        ac.enc_batch(encs, F, h_pcm.data_ptr(), F * N * 2, 16, h_out.data_ptr(), F * NBYTES)   # warm-up (allocates staging)
        barrier()
        t0 = time.perf_counter()
        for _ in range(ksteps):
            ac.enc_batch(encs, F, h_pcm.data_ptr(), F * N * 2, 16, h_out.data_ptr(), F * NBYTES)
        torch.cuda.synchronize()
        e2e_s = max_over_ranks(time.perf_counter() - t0)
        passes += 1 + ksteps
        barrier()
        e2e = {"value": float(S) * F * ksteps * world / e2e_s, "unit": "frames/s", "h2d_bytes_per_step": S * F * N * 2,
               "d2h_bytes_per_step": S * F * NBYTES, "steps": ksteps,
               "note": "pinned host buffers -> lc3plus_enc_batch -> pinned host buffers; copies inside the timed region"}
    stop.set()
    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    # the timed bytes are the reference encoder's bytes: device-resident leg (last timed step) and end-to-end leg (last step)
    v_streams, v_frames = verify_against_reference(pcm, last_out, F, args.warmup + args.steps, "device-resident leg")
    verified = {"verified_streams": v_streams, "verified_frames": v_frames,
                "verified_against": "unmodified ETSI reference encoder (oracle/_ref/refbench) over the same %d passes of the PCM, "
                                    "streams %s of rank 0" % (args.warmup + args.steps, list(VERIFY_STREAMS))}
    if e2e is not None:
        e_streams, e_frames = verify_against_reference(pcm, h_out.numpy(), F, passes, "end-to-end leg")
        e2e["verified_streams"], e2e["verified_frames"] = e_streams, e_frames

    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    hbm_peak, peak_src = (peaks["hbm_gbs"], "measured (MEASURED_PEAKS.json)") if "hbm_gbs" in peaks else (6650.0, "fallback")
    KS = ("stage", "fir", "scalar", "pitch", "mdct", "sns", "tns", "rate", "quant", "back")
    kern = {k: alone[k] / alone["calls"] for k in KS}          # ms per step (S*F channel-frames), kernel alone
    top = max(kern, key=kern.get)
    avg_ms = kern[top]                                          # one launch of the dominant kernel covers the whole step here
    frames_per_launch = S * F
    achieved = ALGO_BYTES_PER_FRAME * frames_per_launch / (avg_ms * 1e-3) / 1e9
    traffic = None
    try:
        # DRAM bytes of the dominant kernel from the committed `ncu --set full` capture (tools/ncu_summary.py), per launch,
        # rescaled linearly if this run's launch covers a different number of channel-frames than the captured one
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        traffic = tj[f"k_{top}"][0]["dram_bytes"] * frames_per_launch / tj["_frames_per_launch"]
    except Exception:
        pass
    sm_clock = summarize_clocks(rows)
    int_peak, int_src = 148 * 4 * 32 * (sm_clock["sm_mhz"] or 1965) * 1e6, "148 SMs x 4 schedulers x 32 lanes x median SM clock"
    try:   # measured on this pool's B200 by tools/int_peak.cu (mixed IMAD+LOP3 stream, 97 % of the issue limit)
        ip = json.load(open(os.path.join(ROOT, "profiles", "int_peak.json")))
        int_peak = ip["mixed_thread_instr_per_s"] * (sm_clock["sm_mhz"] or 1965) * 1e3 / ip["sm_clock_khz_max"]
        int_src = "measured: profiles/int_peak.json (tools/int_peak.cu), scaled to the median SM clock of this run"
    except Exception:
        pass
    roofline = {"bound": "hbm", "kernel": f"k_{top}", "achieved": achieved, "peak": hbm_peak, "unit": "GB/s",
                "frac": achieved / hbm_peak, "traffic": traffic, "peak_source": peak_src,
                "algorithmic_bytes_per_frame": ALGO_BYTES_PER_FRAME, "avg_launch_ms": avg_ms,
                "timing": "CUDA events on the launching stream, each kernel alone on the device (3 untimed-region steps, one "
                          "launch of each kernel per step of %d channel-frames); peak is the burst figure" % (S * F),
                "kernel_ms_alone": {k: round(v, 4) for k, v in kern.items()},
                "kernel_ms_share": {k: v / sum(kern.values()) for k, v in kern.items()},
                "kernel_ms_overlapped_in_timed_region": {k: round(prof[k] / args.steps, 4) for k in KS},
                "int_alu": {"note": "the path is integer-issue bound, not HBM bound (SURVEY.md 8d): reference weighted ops/frame x frames/s "
                                    "over the integer issue peak", "peak_source": int_src,
                            "weighted_ops_per_frame": WOPS_PER_FRAME, "achieved_wops_per_s": value / world * WOPS_PER_FRAME,
                            "peak_thread_instr_per_s": int_peak, "frac": value / world * WOPS_PER_FRAME / int_peak}}
    line = {"metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * dev_s / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "int16/int32 fixed point", "data": "synthetic",
            "config": {"workload": WORKLOAD, "streams_per_gpu": S, "frames_per_step": F, "channel_frames_per_step": S * F * world,
                       "l2": "per-step input 393 MB > 126 MB L2 (no flush needed)"},
            "realtime_channels": value / 100.0, "wall_ms_per_step": 1e3 * wall_s / args.steps,
            "e2e": e2e, "gpu_launches": int(launches), "clocks": sm_clock, "roofline": roofline}
    line.update(verified)
    if world == 1 and not args.no_cpu_baseline and refbench_path() is not None:
        rate, cores, sample = cpu_reference_rate(pcm, F, 12.0)
        line["cpu_baseline"] = {"value": rate, "unit": "frames/s", "cores": cores, "kind": "reference", "sample": sample}
    print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
