#!/usr/bin/env python3
"""Aggregate an ncu report's per-SASS-instruction counters by CUDA source line and by pipeline stage.

    tools/ncu_lines.py <report.ncu-rep> <kernel regex> <mangled name substring> [top N]
    LINES=243,244 tools/ncu_lines.py ...   lists the SASS (executed count, stall samples) of those body lines instead

`ncu --page source --csv` lists SASS instructions in address order with "Instructions Executed" and stall
samples; `nvdisasm -gi -c` on the cubin extracted from liblc3plus_b200.so lists the same instructions with
`//## File "...", line N inlined at "...", line M` chains (nvcc -lineinfo).  The two listings are joined by
position.  Three views: by innermost line (where the instruction was written), by "stage frame" = the outermost
frame of the inline chain that is not the kernel body/wrapper (which stage function called it), and the static
SASS size of each stage (instruction-cache footprint)."""
import csv
import glob
import os
import re
import subprocess
import sys
import tempfile
from collections import defaultdict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# frames that only dispatch: the stage is the first frame below them
WRAPPERS = ("lc3_k_front.cuh", "lc3_k_back.cu", "lc3_k_stream.cu", "lc3_body_frame.cuh", "lc3_k_front_a.cu")
BODIES = ("lc3_pitch_warp.cuh", "lc3_rate_warp.cuh")  # warp bodies: their own lines are the stages


def src_line(cache, f, ln):
    path = os.path.join(ROOT, "audio_codec_b200", "csrc", f)
    if path not in cache:
        cache[path] = open(path).read().splitlines() if os.path.exists(path) else []
    L = cache[path]
    return L[ln - 1].strip()[:84] if 0 < ln <= len(L) else ""


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
            dis = subprocess.run(["nvdisasm", "-gi", "-c", cub], capture_output=True, text=True).stdout.splitlines()
            if any(l.startswith(".text.") and mangled in l for l in dis):
                break
    start = next(i for i, l in enumerate(dis) if l.startswith(".text.") and mangled in l)
    chains = []          # per instruction: tuple of (file, line) frames, innermost first
    cur, fresh = [], True
    for l in dis[start + 1:]:
        if l.startswith("\t.section") or l.startswith(".text."):
            break
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            if fresh:
                cur, fresh = [], False
            cur.append((os.path.basename(m.group(1)), int(m.group(2))))
            continue
        if re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(.*?);", l):
            chains.append(tuple(cur) if cur else (("?", 0),))
            fresh = True
    n = min(len(chains), len(sass))
    if len(chains) != len(sass):
        print(f"warning: {len(sass)} profiled vs {len(chains)} disassembled instructions", file=sys.stderr)

    def stage_of(ch):
        # the call site in the kernel body: innermost frame that lies in a body/wrapper file
        for fr in ch:
            if fr[0] in WRAPPERS or fr[0] in BODIES:
                return fr
        return ch[-1]

    def below_stage(ch):
        # the line of the stage function itself: the frame just inside the call site
        for i, fr in enumerate(ch):
            if fr[0] in WRAPPERS or fr[0] in BODIES:
                return ch[i - 1] if i > 0 else fr
        return ch[0]

    if os.environ.get("LINES"):  # dump the SASS (with executed counts) whose call site in the body is one of these lines
        want = {int(v) for v in os.environ["LINES"].split(",")}
        for k in range(n):
            if stage_of(chains[k])[1] in want or (os.environ.get("LINES_FILE") and any(fr[0] == os.environ["LINES_FILE"] and fr[1] in want for fr in chains[k])):
                print(f"{k:6d} {sass[k][1]:12d} {sass[k][2]:6d}  L{stage_of(chains[k])[1]:<4d} {sass[k][0]}")
        return
    tot_i = sum(s[1] for s in sass[:n])
    tot_s = sum(s[2] for s in sass[:n])
    print(f"total warp instructions {tot_i}, stall samples {tot_s}, SASS instructions {n}")
    inner, stage, stagefile = defaultdict(lambda: [0, 0, 0]), defaultdict(lambda: [0, 0, 0]), defaultdict(lambda: [0, 0, 0])
    for k in range(n):
        for d, key in ((inner, chains[k][0]), (stage, stage_of(chains[k])), (stagefile, below_stage(chains[k]))):
            a = d[key]
            a[0] += sass[k][1]; a[1] += sass[k][2]; a[2] += 1
    cache = {}
    pct = lambda a: f"inst {100 * a[0] / max(tot_i, 1):5.1f}%  samp {100 * a[1] / max(tot_s, 1):5.1f}%  sass {a[2]:5d}"
    print("by call site in the kernel body (everything inlined below it included; exact, from the inline chain):")
    for (f, ln), a in sorted(stage.items(), key=lambda x: -x[1][0])[:top]:
        print(f"  {f}:{ln:<4d} {pct(a)} | {src_line(cache, f, ln)}")
    print("by line of the stage function called from the body (its callees included):")
    for (f, ln), a in sorted(stagefile.items(), key=lambda x: -x[1][0])[:top]:
        print(f"  {f}:{ln:<4d} {pct(a)} | {src_line(cache, f, ln)}")
    print("top innermost lines:")
    for (f, ln), a in sorted(inner.items(), key=lambda x: -x[1][0])[:top]:
        print(f"  {f}:{ln:<4d} {pct(a)} | {src_line(cache, f, ln)}")


if __name__ == "__main__":
    main()
