#!/usr/bin/env python3
"""Run the harness binaries of baseline/build.py on the GPU box and summarise their own log lines.
    python baseline/run.py [out_dir]
<sample>_b200 = unmodified sample driver + libb200prim.so (must exit 0: the sample's own gold check passed);
<sample>_ref  = the reference's own kernels recompiled for sm_100a (the baseline to beat).
Throughput figures are the samples' OWN measurements (host wall clock around their timed loops)."""
import json
import os
import re
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
BIN = os.path.join(HERE, "_ref")
RUNS = [
    ("histogram", []),
    ("reduction", ["--type=int", "--n=268435456"]),
    ("reduction", ["--type=float", "--n=268435456"]),
    ("reduction", ["--type=double", "--n=134217728"]),
    ("scan", []),
    ("mergeSort", []),
    ("sortingNetworks", []),
    ("shfl_scan", []),
    ("warpAggregatedAtomicsCG", []),
]


def main(out_dir):
    os.makedirs(out_dir, exist_ok=True)
    summary = []
    for name, args in RUNS:
        for flavour in ("b200", "ref"):
            exe = os.path.join(BIN, "%s_%s" % (name, flavour))
            if not os.path.exists(exe):
                continue
            r = subprocess.run([exe] + args, capture_output=True, text=True, timeout=600)
            tag = "%s_%s%s" % (name, flavour, "".join("_" + a.strip("-").replace("=", "") for a in args))
            with open(os.path.join(out_dir, "baseline_%s.log" % tag), "w") as f:
                f.write(r.stdout + "\n--- stderr ---\n" + r.stderr)
            lines = [l.strip() for l in r.stdout.splitlines()
                     if re.search(r"Throughput|Time:|Time \(|MegaElements|CheckSum|stable|Test passed|Test failed|CPU result|GPU result", l)]
            summary.append({"run": tag, "exit_code": r.returncode, "lines": lines})
            print(tag, "exit", r.returncode)
            for l in lines:
                print("    ", l)
    with open(os.path.join(out_dir, "baseline_summary.json"), "w") as f:
        json.dump(summary, f, indent=1)
    bad = [s["run"] for s in summary if s["run"].split("_")[1] == "b200" and s["exit_code"] != 0]
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(HERE), "gpurun_out")))
