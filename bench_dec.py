#!/usr/bin/env python3
"""Decoder bench -- companion of bench.py, which stays the headline (encoder) bench.

    python bench_dec.py [--steps K] [--warmup W] [--streams S] [--frames F] [--config c1|c3]

A step decodes F consecutive frames of each of S streams (BASELINE.json geometry of the configuration) with ONE call of
lc3plus_dec_batch; consecutive steps continue the streams (state carried), decoding the same payloads again.  Payloads come from
the GPU encoder (bit-exact with the reference encoder, tests/test_gpu_parity.py).  Prints ONE JSON line:

  value     decoded channel-frames/s with payloads and PCM resident in HBM (device span of the calls, CUDA events in the library)
  e2e       the same through the C ABI with PINNED HOST buffers, host<->device copies inside the timed region (wall clock
            around the synchronous call)
  verified  PCM of the last timed step of both legs == the unmodified reference decoder's (oracle/_ref/refbench_dec, same
            payloads, same number of passes) on a sample of streams
  roofline  integer-issue fraction: reference decoder's weighted ops per frame (profiles/wmops_dec.json; the count includes
            the concealment bookkeeping it runs on good frames) x frames/s over the measured integer issue peak; kernel ms of
            the first chunk of every call
  cpu_baseline  the reference decoder on all host cores (oracle/_ref/refbench_dec), bounded sample
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
CONFIGS = {"c1": (48000, 100, 128000, 4096, 1000), "c3": (16000, 50, 32000, 12500, 1000)}


def refdec(payloads, fs, dms, nb, nf, threads, out=None, repeat=1):
    exe = os.path.join(ROOT, "oracle", "_ref", "refbench_dec")
    with tempfile.TemporaryDirectory() as d:
        p = os.path.join(d, "in.bin")
        payloads.tofile(p)
        o = os.path.join(d, "out.pcm") if out else "-"
        r = subprocess.run([exe, p, str(fs), str(dms), str(nb), str(payloads.shape[0]), str(nf), str(threads), o, str(repeat)],
                           capture_output=True, text=True, check=True).stdout.split()
        pcm = np.fromfile(o, dtype=np.int16).reshape(payloads.shape[0], -1) if out else None
    return float(r[2]), float(r[1]), int(r[4]), pcm


def quick_leg(S=4096, nf=100, fs=48000, dms=100, br=128000, steps=5, warmup=2):
    """A short decoder leg for bench.py's JSON line (key "decoder"): S streams x nf frames per step, device-resident, after the
    encoder's own run; four streams of a fresh one-call decode are checked against the decoder oracle (oracle/lc3o_dec.c)."""
    import torch
    import audio_codec_b200 as ac
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import lc3ref
    from audio_codec_b200.synth import make_streams_fast
    N = fs // 100 * dms // 100
    pcm = make_streams_fast(S, nf * N, fs, N, 16)
    encs = ac.EncoderArray([ac.Encoder(fs, 1, dms, br) for _ in range(S)])
    nb = encs.encoders[0].num_bytes
    d_pcm = torch.from_numpy(pcm).cuda()
    d_bytes = torch.zeros((S, nf * nb), dtype=torch.uint8, device="cuda")
    ac.enc_batch(encs, nf, d_pcm.data_ptr(), nf * N * 2, 16, d_bytes.data_ptr(), nf * nb)
    for e in encs.encoders:
        e.close()
    nbs = np.full(S, nb, np.int32)
    d_out = torch.zeros((S, nf * N), dtype=torch.int16, device="cuda")
    decs = ac.DecoderArray([ac.Decoder(fs, 1, dms) for _ in range(S)])
    assert ac.dec_batch(decs, nf, d_bytes.data_ptr(), nf * nb, nbs, d_out.data_ptr(), nf * N * 2, 16) == 0
    idx = [0, 7, S // 2, S - 1]
    got = d_out[idx].cpu().numpy()
    payloads = d_bytes[idx].cpu().numpy().reshape(len(idx), nf, nb)
    for k in range(len(idx)):
        ref, bad = lc3ref.OracleDec(fs, 1, dms).decode(payloads[k], 16)
        assert bad == 0 and np.array_equal(ref[0], got[k]), "decoder leg differs from the decoder oracle"
    for _ in range(warmup):
        ac.dec_batch(decs, nf, d_bytes.data_ptr(), nf * nb, nbs, d_out.data_ptr(), nf * N * 2, 16)
    ac.dec_profile_enable(True)
    n0 = ac.dec_kernel_launches()
    for _ in range(steps):
        ac.dec_batch(decs, nf, d_bytes.data_ptr(), nf * nb, nbs, d_out.data_ptr(), nf * N * 2, 16)
    span = ac.dec_profile_read()["span"] / 1e3
    launches = ac.dec_kernel_launches() - n0
    ac.dec_profile_enable(False)
    for d in decs.decoders:
        d.close()
    return {"metric": "decoded channel-frames per second, device-resident (SURVEY.md 8 row f4; full bench: bench_dec.py)", "value": S * nf * steps / span,
            "unit": "frames/s", "steps": steps, "warmup": warmup, "ms_per_step": span * 1e3 / steps, "gpu_launches": launches,
            "verified_frames": len(idx) * nf, "workload": f"{S} streams x {nf} frames/step, {fs} Hz, {dms / 10} ms, {nb} bytes/frame"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--config", default="c1", choices=sorted(CONFIGS))
    ap.add_argument("--streams", type=int, default=0)
    ap.add_argument("--frames", type=int, default=0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--gpus", type=int, default=1, help="under torchrun: one process per GPU, each decoding its own S streams (weak scaling, no collective)")
    args = ap.parse_args()
    import torch
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(v):
        if dist is None:
            return v
        t = torch.tensor([v], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())
    import audio_codec_b200 as ac
    import bench
    from audio_codec_b200.synth import make_streams_fast
    fs, dms, br, S, nf = CONFIGS[args.config]
    S, nf = args.streams or S, args.frames or nf
    N = fs // 100 * dms // 100
    pcm = make_streams_fast(S, nf * N, fs, N, 16)
    encs = ac.EncoderArray([ac.Encoder(fs, 1, dms, br) for _ in range(S)])
    nb = encs.encoders[0].num_bytes
    d_pcm = torch.from_numpy(pcm).cuda()
    d_bytes = torch.zeros((S, nf * nb), dtype=torch.uint8, device="cuda")
    ac.enc_batch(encs, nf, d_pcm.data_ptr(), nf * N * 2, 16, d_bytes.data_ptr(), nf * nb)
    del d_pcm, pcm
    for e in encs.encoders:
        e.close()
    nbs = np.full(S, nb, np.int32)
    d_out = torch.zeros((S, nf * N), dtype=torch.int16, device="cuda")
    passes = args.warmup + args.steps

    # ---- device-resident leg
    decs = ac.DecoderArray([ac.Decoder(fs, 1, dms) for _ in range(S)])
    for _ in range(args.warmup):
        ac.dec_batch(decs, nf, d_bytes.data_ptr(), nf * nb, nbs, d_out.data_ptr(), nf * N * 2, 16)
    rows, stop = [], threading.Event()
    th = threading.Thread(target=bench.clock_sampler, args=(stop, rows, torch.cuda.current_device()))
    th.start()
    ac.dec_profile_enable(True)
    n0 = ac.dec_kernel_launches()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        rc = ac.dec_batch(decs, nf, d_bytes.data_ptr(), nf * nb, nbs, d_out.data_ptr(), nf * N * 2, 16)
        assert rc == 0
    barrier()
    wall = time.perf_counter() - t0
    prof = ac.dec_profile_read()
    launches = ac.dec_kernel_launches() - n0
    ac.dec_profile_enable(False)
    stop.set(); th.join()
    clocks = bench.summarize_clocks(rows)
    span_s = max_over_ranks(prof["span"] / 1e3)
    value = world * S * nf * args.steps / span_s
    idx = sorted(set([0, 7, S // 3, S // 2, S - 1] + list(range(8, min(S, 24)))))
    out_dev = d_out[idx].cpu().numpy()   # the rows that are checked against the reference below
    del d_out
    for d in decs.decoders:
        d.close()

    # ---- end-to-end leg: pinned host payloads and PCM
    t_in = torch.empty((S, nf * nb), dtype=torch.uint8, pin_memory=True)
    t_in.copy_(d_bytes)
    t_out = torch.empty((S, nf * N), dtype=torch.int16, pin_memory=True)
    decs = ac.DecoderArray([ac.Decoder(fs, 1, dms) for _ in range(S)])
    for _ in range(args.warmup):
        ac.dec_batch(decs, nf, t_in.numpy(), nf * nb, nbs, t_out.numpy(), nf * N * 2, 16)
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ac.dec_batch(decs, nf, t_in.numpy(), nf * nb, nbs, t_out.numpy(), nf * N * 2, 16)
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    e2e = world * S * nf * args.steps / e2e_s
    out_e2e = t_out.numpy()[idx].copy()

    # ---- both legs against the reference decoder, same payloads, same number of passes
    payloads = t_in.numpy().reshape(S, nf, nb)
    verified = None
    have_ref = os.path.exists(os.path.join(ROOT, "oracle", "_ref", "refbench_dec"))
    if have_ref:
        _, _, bad, ref = refdec(np.ascontiguousarray(payloads[idx]), fs, dms, nb, nf, min(len(idx), os.cpu_count() or 1), out=True, repeat=passes)
        assert bad == 0
        assert np.array_equal(ref, out_dev), "device-resident leg differs from the reference decoder"
        assert np.array_equal(ref, out_e2e), "end-to-end leg differs from the reference decoder"
        verified = len(idx) * nf

    cpu = None
    if have_ref and not args.no_cpu_baseline and world == 1:
        cores = os.cpu_count() or 1
        ns = min(S, max(cores * 2, 32))
        rate, secs, _, _ = refdec(np.ascontiguousarray(payloads[:ns, :200]), fs, dms, nb, 200, cores)
        rep = int(max(1, min(200, round(12.0 / max(secs, 1e-3)))))
        rate, secs, bad, _ = refdec(np.ascontiguousarray(payloads[:ns, :200]), fs, dms, nb, 200, cores, repeat=rep)
        cpu = {"value": rate, "unit": "decoded channel-frames/s", "cores": cores, "kind": "reference",
               "sample": f"{ns} of the workload's streams x {200 * rep} frames ({secs:.1f} s) on {cores} threads, gcc -O2, reference decoder with its concealment bookkeeping"}

    roof = {"kernel_ms_first_chunk": {k: round(v / args.steps, 4) for k, v in prof.items() if k != "span"}}
    try:
        w = json.load(open(os.path.join(ROOT, "profiles", "wmops_dec.json")))[args.config]
        ip = json.load(open(os.path.join(ROOT, "profiles", "int_peak.json")))
        int_peak = ip["mixed_thread_instr_per_s"] * (clocks["sm_mhz"] or 1965) * 1e3 / ip["sm_clock_khz_max"]
        wops = w["weighted_ops_per_channel_frame"]
        roof.update({"bound": "int_alu", "weighted_ops_per_channel_frame_reference": wops,
                     "note": "reference decoder weighted ops per frame on this stimulus (profiles/wmops_dec_%s.txt; includes its concealment "
                             "bookkeeping on good frames, about 10 %%, which this library does not run)" % args.config,
                     "achieved": value * wops, "peak": int_peak, "unit": "weighted op/s vs thread instr/s", "frac": value * wops / int_peak,
                     "hbm_algorithmic_GBps": value * (nb + N * 2) / 1e9})
    except (OSError, KeyError):
        pass
    line = {"metric": "decoded LC3plus channel-frames per second (bit-exact with the reference decoder, error-free streams)", "value": value,
            "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": span_s * 1e3 / args.steps,
            "wall_ms_per_step": wall * 1e3 / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "int16/int32 fixed point",
            "data": "synthetic (audio_codec_b200.synth) encoded by the GPU encoder", "path": "decoder",
            "config": {"workload": f"{args.config} geometry, decode: {S} streams per GPU x {nf} frames/step, {fs} Hz, {dms / 10} ms, {nb} bytes/frame",
                       "inputs": "larger than L2 (payloads %.0f MB, PCM %.0f MB per step)" % (S * nf * nb / 1e6, S * nf * N * 2 / 1e6)},
            "e2e": {"value": e2e, "unit": "frames/s", "h2d_bytes_per_step": S * nf * nb, "d2h_bytes_per_step": S * nf * N * 2,
                    "pcie_GBps": world * S * nf * args.steps * (nb + N * 2) / e2e_s / 1e9},
            "gpu_launches": launches, "verified_frames": verified, "roofline": roof, "cpu_baseline": cpu, "clocks": clocks}
    if rank == 0:
        print(json.dumps(line))
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
