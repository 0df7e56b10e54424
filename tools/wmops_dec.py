#!/usr/bin/env python
"""Weighted-operation counts of the REFERENCE DECODER on the benchmark stimulus: runs oracle/_ref/LC3plus_wmops (the unmodified
reference built with its complexity counters on) as encoder and then as decoder over streams 0..7 of audio_codec_b200.synth and
writes profiles/wmops_dec_<cfg>.txt + profiles/wmops_dec.json (weighted ops per decoded CHANNEL-frame; bench_dec.py reads it).
Needs /root/reference."""
import json
import os
import re
import subprocess
import sys
import tempfile
import wave

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from audio_codec_b200.synth import make_streams  # noqa: E402

CFG = {"c1": (48000, 1, 100, 128000), "c3": (16000, 1, 50, 32000), "c4": (48000, 2, 25, 256000)}
SECONDS = 6
EXE = os.path.join(ROOT, "oracle", "_ref", "LC3plus_wmops")


def run(fs, ch, dms, br, first):
    N = fs // 100 * dms // 100
    x = make_streams(ch, SECONDS * fs, fs, N, first_stream=first)
    with tempfile.TemporaryDirectory() as d:
        with wave.open(os.path.join(d, "in.wav"), "wb") as w:
            w.setnchannels(ch); w.setsampwidth(2); w.setframerate(fs)
            w.writeframes(np.ascontiguousarray(x.T).tobytes())
        subprocess.run([EXE, "-E", "-q", "-frame_ms", str(dms / 10.0), os.path.join(d, "in.wav"), os.path.join(d, "out.lc3"), str(br)],
                       capture_output=True, text=True, check=True)
        o = subprocess.run([EXE, "-D", "-q", os.path.join(d, "out.lc3"), os.path.join(d, "out.wav")], capture_output=True, text=True, check=True).stdout
    i = o.index("Weighted MOPS Analysis")
    table = o[i:].rstrip()
    m = re.search(r"^total\s+[\d.]+\s+([\d.]+)\s+([\d.]+)\s+([\d.]+)\s+([\d.]+)", table, re.M)
    return table, float(m.group(3)), float(m.group(2))


def main():
    out = {}
    for name, (fs, ch, dms, br) in CFG.items():
        rows, table1 = [], None
        for s in range(8):
            table, avg, worst = run(fs, ch, dms, br, s * ch)
            rows.append((s, avg, worst))
            if s == 1:
                table1 = table
        fps = 1000.0 / (dms / 10.0)
        avg = float(np.mean([r[1] for r in rows]))
        per_cf = avg * 1e6 / fps / ch
        out[name] = {"fs": fs, "channels": ch, "dms": dms, "bitrate": br, "wmops_avg": avg, "wmops_worst": max(r[2] for r in rows),
                     "weighted_ops_per_channel_frame": per_cf, "streams": 8, "seconds_per_stream": SECONDS}
        with open(os.path.join(ROOT, "profiles", f"wmops_dec_{name}.txt"), "w") as f:
            f.write(f"reference DECODER built with WMOPS=1 (fx/makefile:8), {fs} Hz, {ch} ch, {dms / 10} ms, {br} bps, synth.make_streams streams 0..7, {SECONDS} s each\n")
            f.write("stream  avg WMOPS  worst-case WMOPS\n")
            for s, a, w in rows:
                f.write(f"{s:6d}  {a:9.3f}  {w:9.3f}\n")
            f.write(f"mean of the averages: {avg:.3f} WMOPS = {per_cf:.0f} weighted ops per channel-frame (includes the concealment bookkeeping the reference runs on good frames)\n\nper-stage table of stream 1:\n{table1}\n")
        print(name, round(avg, 3), round(per_cf))
    json.dump(out, open(os.path.join(ROOT, "profiles", "wmops_dec.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
