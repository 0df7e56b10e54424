#!/usr/bin/env python3
"""Aggregate an ncu report's per-SASS-instruction counters by CUDA source line.

    tools/ncu_lines.py <report.ncu-rep> <kernel regex> <mangled name substring> [top N]

`ncu --page source --csv` lists SASS instructions in address order with "Instructions Executed" and stall
samples; `nvdisasm -g -c` on the cubin extracted from liblc3plus_b200.so lists the same instructions with
`//## File "...", line N` markers (nvcc -lineinfo).  The two listings are joined by position."""
import csv
import glob
import os
import re
import subprocess
import sys
import tempfile
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    rep, kre, mangled = sys.argv[1], sys.argv[2], sys.argv[3]
    top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", f"regex:{kre}"],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    cols = rows[hdr]
    ci, cs, csrc = cols.index("Instructions Executed"), cols.index("# Samples"), cols.index("Source")
    sass = []
    for r in rows[hdr + 1:]:
        if len(r) <= ci or not r[0].startswith("0x"):
            break  # first kernel instance only
        sass.append((r[csrc].strip(), int(float(r[ci] or 0)), int(float(r[cs] or 0))))
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", os.path.join(ROOT, "audio_codec_b200", "liblc3plus_b200.so")], cwd=tmp,
                   capture_output=True)
    dis = []
    for cub in glob.glob(os.path.join(tmp, "*.cubin")):
        if mangled.encode() in open(cub, "rb").read():
            dis = subprocess.run(["nvdisasm", "-g", "-c", cub], capture_output=True, text=True).stdout.splitlines()
            if any(l.startswith(".text.") and mangled in l for l in dis):
                break
    start = next(i for i, l in enumerate(dis) if l.startswith(".text.") and mangled in l)
    lines = []
    cur = ("?", 0)
    stack = ""
    for l in dis[start + 1:]:
        if l.startswith("\t.section") or l.startswith(".text."):
            break
        m = re.search(r'//## File "([^"]+)", line (\d+)(.*)', l)
        if m:
            cur = (os.path.basename(m.group(1)), int(m.group(2)))
            stack = m.group(3)
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(.*?);", l)
        if m:
            lines.append(cur)
    n = min(len(lines), len(sass))
    if len(lines) != len(sass):
        print(f"warning: {len(sass)} profiled vs {len(lines)} disassembled instructions", file=sys.stderr)
    agg = defaultdict(lambda: [0, 0, 0])
    tot_i = tot_s = 0
    for k in range(n):
        a = agg[lines[k]]
        a[0] += sass[k][1]; a[1] += sass[k][2]; a[2] += 1
        tot_i += sass[k][1]; tot_s += sass[k][2]
    print(f"total warp instructions {tot_i}, stall samples {tot_s}, SASS instructions {n}")
    byfile = defaultdict(lambda: [0, 0])
    for (f, ln), a in agg.items():
        byfile[f][0] += a[0]; byfile[f][1] += a[1]
    for f, a in sorted(byfile.items(), key=lambda x: -x[1][0]):
        print(f"  {f:28s} inst {100 * a[0] / max(tot_i, 1):5.1f}%  samples {100 * a[1] / max(tot_s, 1):5.1f}%")
    # primitives (lc3_basop/lc3_math/intrinsics) carry no inline stack in the line table: attribute them to the
    # nearest preceding line of stage code
    prim = ("lc3_basop.cuh", "lc3_math.cuh", "lc3_warp.cuh", "sm_30_intrinsics.hpp", "sm_80_rt.hpp")
    owner, bycaller, bycl = ("?", 0), defaultdict(lambda: [0, 0]), defaultdict(lambda: [0, 0])
    for k in range(n):
        if lines[k][0] not in prim:
            owner = lines[k]
        bycaller[owner[0]][0] += sass[k][1]; bycaller[owner[0]][1] += sass[k][2]
        bycl[owner][0] += sass[k][1]; bycl[owner][1] += sass[k][2]
    print("by stage file (primitives attributed to the nearest preceding stage line):")
    for f, a in sorted(bycaller.items(), key=lambda x: -x[1][0]):
        print(f"  {f:28s} inst {100 * a[0] / max(tot_i, 1):5.1f}%  samples {100 * a[1] / max(tot_s, 1):5.1f}%")
    print("top stage lines incl. their primitives:")
    for (f, ln), a in sorted(bycl.items(), key=lambda x: -x[1][0])[:top // 2]:
        path = os.path.join(ROOT, "audio_codec_b200", "csrc", f)
        txt = ""
        if os.path.exists(path):
            L = open(path).read().splitlines()
            txt = L[ln - 1].strip()[:80] if 0 < ln <= len(L) else ""
        print(f"  {f}:{ln:<4d} inst {100 * a[0] / max(tot_i, 1):5.1f}%  samp {100 * a[1] / max(tot_s, 1):5.1f}% | {txt}")
    print("top lines by executed instructions:")
    src_cache = {}
    for (f, ln), a in sorted(agg.items(), key=lambda x: -x[1][0])[:top]:
        path = os.path.join(ROOT, "audio_codec_b200", "csrc", f)
        if path not in src_cache:
            src_cache[path] = open(path).read().splitlines() if os.path.exists(path) else []
        text = src_cache[path][ln - 1].strip()[:90] if 0 < ln <= len(src_cache[path]) else ""
        print(f"  {f}:{ln:<4d} inst {100 * a[0] / max(tot_i, 1):5.1f}%  samp {100 * a[1] / max(tot_s, 1):5.1f}%  sass {a[2]:4d} | {text}")


if __name__ == "__main__":
    main()
