#!/usr/bin/env python3
"""Summarise an ncu report (read here, no GPU): headline metrics of each captured launch and the warp-stall samples of
the first launch grouped by the kernel's barrier-delimited regions, plus the instructions that collect the most samples.
    python profiles/ncu_regions.py gpurun_out/<name>.ncu-rep [top_n]"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
top_n = int(sys.argv[2]) if len(sys.argv) > 2 else 25
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, data = rows[0], rows[2:]
want = ["Kernel Name", "gpu__time_duration.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "l1tex__data_pipe_lsu_wavefronts.sum", "smsp__inst_executed.sum", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size"]
for h in want:
    if h in hdr:
        i = hdr.index(h)
        print("%-70s %s" % (h, [r[i][:60] for r in data]))
for i, h in enumerate(hdr):
    if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("per_issue_active.ratio"):
        v = [float(r[i] or 0) for r in data]
        if max(v) > 0.3:
            print("%-70s %s" % (h.replace("smsp__average_warps_issue_stalled_", "stall/issue: ").replace(
                "_per_issue_active.ratio", ""), ["%.2f" % x for x in v]))

src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True,
                     text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hdr = rows[1]
idx = {h: i for i, h in enumerate(hdr)}
body = []
for r in rows[2:]:
    if r and r[0] == "Kernel Name":
        break
    if len(r) == len(hdr):
        body.append(r)
tot = sum(int(r[idx["# Samples"]] or 0) for r in body) or 1
stall = [c for c in hdr if c.startswith("stall_") and "(" not in c]
print("\nfirst launch: %d instructions, %d samples; regions end at BAR.SYNC / EXIT" % (len(body), tot))
cur, n0 = None, 0
for i, r in enumerate(body):
    if cur is None:
        cur, n0 = {"s": 0, "exec": 0, **{c: 0 for c in stall}}, i
    cur["s"] += int(r[idx["# Samples"]] or 0)
    cur["exec"] += int(r[idx["Instructions Executed"]] or 0)
    for c in stall:
        cur[c] += int(r[idx[c]] or 0)
    text = r[idx["Source"]]
    if "BAR.SYNC" in text or "EXIT" in text or i == len(body) - 1:
        if cur["s"] > tot * 0.003:
            top = sorted(((cur[c], c[6:]) for c in stall), reverse=True)[:5]
            print("  instr %4d-%4d  %5.1f%% of samples  %10d warp-instr  | %s" % (
                n0, i, 100 * cur["s"] / tot, cur["exec"], " ".join("%s=%.1f%%" % (n, 100 * v / tot) for v, n in top if v)))
        cur = None
print("\ntop instructions by samples:")
order = sorted(range(len(body)), key=lambda i: -int(body[i][idx["# Samples"]] or 0))[:top_n]
for i in sorted(order):
    r = body[i]
    st = sorted(((int(r[idx[c]] or 0), c[6:]) for c in stall), reverse=True)[:2]
    print("  %4d %-58s %5.1f%%  %s" % (i, r[idx["Source"]].strip()[:58], 100 * int(r[idx["# Samples"]]) / tot,
                                      " ".join("%s=%d" % (n, v) for v, n in st if v)))
