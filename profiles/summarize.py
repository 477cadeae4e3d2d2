#!/usr/bin/env python3
"""ncu report -> small CSV of the metrics the roofline claims rest on.
    python profiles/summarize.py gpurun_out/prof.ncu-rep profiles/rNN_summary.csv
Also refreshes profiles/traffic.json (dram bytes per launch of each kernel family, read by bench.py)."""
import csv
import json
import os
import subprocess
import sys

WANT = ["Kernel Name", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__cycles_active.avg", "sm__cycles_elapsed.max"]
SCALE = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}


def reduce_source_hash():
    """sha256 over the sources reduce_kernel is compiled from: ties profiles/traffic.json to the kernel that is timed."""
    import hashlib
    csrc = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "cuda_samples_b200", "csrc")
    h = hashlib.sha256()
    for name in ("reduce.cu", "common.cuh", "p2p.cuh", "internal.h"):
        with open(os.path.join(csrc, name), "rb") as f:
            h.update(f.read())
    return h.hexdigest()


def main(rep, out):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    cols = [c for c in WANT if c in idx]
    traffic = {}
    with open(out, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(cols)
        w.writerow([units[idx[c]] for c in cols])
        for r in rows[2:]:
            w.writerow([r[idx[c]] for c in cols])
            name = r[idx["Kernel Name"]].split("(")[0].replace("void ", "").strip()
            rd = float(r[idx["dram__bytes_read.sum"]]) * SCALE[units[idx["dram__bytes_read.sum"]]]
            wr = float(r[idx["dram__bytes_write.sum"]]) * SCALE[units[idx["dram__bytes_write.sum"]]]
            traffic.setdefault(name + "@grid" + r[idx["launch__grid_size"]], []).append(rd + wr)
    summary = {k: sum(v) / len(v) for k, v in traffic.items()}
    red = [v for k, v in summary.items() if "reduce_kernel" in k]
    if red:
        summary["reduce_kernel"] = max(red)
    summary["_reduce_source_sha256"] = reduce_source_hash()  # bench.py refuses the figure when the kernel has changed
    summary["_source"] = os.path.basename(rep) + ": ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum per launch"
    tj = os.path.join(os.path.dirname(os.path.abspath(out)), "traffic.json")
    try:  # a capture restricted with -k keeps the other kernels' earlier figures
        with open(tj) as f:
            for k, v in json.load(f).items():
                if k not in summary and not k.startswith("_"):
                    summary[k] = v
    except Exception:
        pass
    with open(tj, "w") as f:
        json.dump(summary, f, indent=1, sort_keys=True)
    print("wrote", out, tj)


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
