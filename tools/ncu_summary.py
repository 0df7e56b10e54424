#!/usr/bin/env python
"""Summarise an `ncu --set full` report into the few numbers DESIGN.md / bench.py quote.

    python tools/ncu_summary.py gpurun_out/prof.ncu-rep profiles/r1_full_summary.md [profiles/traffic.json [channel-frames per launch]]

Per kernel launch: duration, DRAM bytes read+written (-> traffic.json, which bench.py reads for
`roofline.traffic`), executed warp instructions, IPC, issue-slot and pipe utilisation, registers,
occupancy, average active threads per executed instruction, and the three largest warp-stall reasons.
"""
import csv
import json
import subprocess
import sys

UNIT = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12,
        "ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}  # times -> ms, bytes -> B


def load(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    return rows[0], rows[1], rows[2:]


def main():
    rep, out_md = sys.argv[1], sys.argv[2]
    out_json = sys.argv[3] if len(sys.argv) > 3 else None
    H, U, rows = load(rep)
    col = {h: i for i, h in enumerate(H)}

    def val(r, name, scale=True):
        i = col.get(name)
        if i is None or r[i] == "":
            return None
        v = float(r[i].replace(",", ""))
        return v * UNIT.get(U[i], 1) if scale else v

    stall_cols = [h for h in H if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
    if not stall_cols:
        stall_cols = [h for h in H if h.startswith("smsp__average_warp_latency_issue_stalled_")]
    lines = ["| kernel | ms | DRAM rd MB | DRAM wr MB | DRAM GB/s | warp inst (M) | IPC/SM | issue % | alu % | fma % | lsu % | regs | occ % (theor.) | thr/inst | top stalls |",
             "|---|---|---|---|---|---|---|---|---|---|---|---|---|---|---|"]
    traffic = {}
    for r in rows:
        name = r[col["Kernel Name"]]
        short = name.split("(")[0].replace("void ", "")
        ms = val(r, "gpu__time_duration.sum")
        rd, wr = val(r, "dram__bytes_read.sum"), val(r, "dram__bytes_write.sum")
        inst = val(r, "smsp__inst_executed.sum", False) or val(r, "sm__inst_executed.sum", False)
        ipc = val(r, "sm__inst_executed.avg.per_cycle_active", False)
        issue = val(r, "sm__inst_issued.avg.pct_of_peak_sustained_active", False) or val(r, "smsp__issue_active.avg.pct_of_peak_sustained_active", False)
        alu = val(r, "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", False)
        fma = val(r, "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", False)
        lsu = val(r, "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", False)
        regs = val(r, "launch__registers_per_thread", False)
        occ = val(r, "sm__warps_active.avg.pct_of_peak_sustained_active", False)
        occ_t = val(r, "sm__maximum_warps_per_active_cycle_pct", False)
        tpi = val(r, "smsp__thread_inst_executed_per_inst_executed.ratio", False)
        stalls = sorted(((val(r, c, False) or 0.0, c) for c in stall_cols), reverse=True)[:3]
        st = ", ".join("%s %.1f" % (c.split("issue_stalled_")[1].split("_per_")[0].replace(".ratio", ""), v) for v, c in stalls)

        def f(x, fmt="%.1f"):
            return "-" if x is None else fmt % x
        lines.append("| %s | %.3f | %s | %s | %s | %s | %s | %s | %s | %s | %s | %s | %s (%s) | %s | %s |" % (
            short, ms, f(rd and rd / 1e6), f(wr and wr / 1e6), f((rd + wr) / 1e9 / (ms * 1e-3)) if rd is not None else "-",
            f(inst and inst / 1e6), f(ipc, "%.2f"), f(issue), f(alu), f(fma), f(lsu), f(regs, "%d"), f(occ), f(occ_t), f(tpi), st))
        key = short.split("<")[0]
        if rd is not None:
            traffic.setdefault(key, []).append({"kernel": short, "ms_under_ncu": ms, "dram_read_bytes": rd, "dram_write_bytes": wr,
                                                "dram_bytes": rd + wr})
    open(out_md, "w").write("Source: `%s` (ncu --set full --clock-control none; per-launch values, cold cache, serialised)\n\n" % rep
                            + "\n".join(lines) + "\n")
    if out_json:
        traffic["_frames_per_launch"] = int(sys.argv[4]) if len(sys.argv) > 4 else 409600
        json.dump(traffic, open(out_json, "w"), indent=1)
    print("\n".join(lines))


if __name__ == "__main__":
    main()
