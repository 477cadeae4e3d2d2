#!/usr/bin/env python3
"""Static SASS opcode listing of every kernel in libb200prim.so (cuobjdump -sass; no GPU needed): the mnemonics that
prove which hardware paths the kernels use (UBLKCP = cp.async.bulk / 1-D TMA, SYNCS = mbarrier, UCGABAR = cluster
barrier, REDUX = redux.sync, ATOMS / RED / ATOMG = shared / global atomics, VOTE = ballot ranking, ACQBULK / PREEXIT =
programmatic dependent launch, LDG.E.128 / STG.E.128 = 16-byte accesses, MEMBAR / FENCE / ERRBAR = ordering).
    python profiles/sass_static.py [lib.so] > profiles/r02_sass_opcodes.txt"""
import collections
import os
import re
import subprocess
import sys

lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))),
                                                         "cuda_samples_b200", "libb200prim.so")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
demangle = lambda n: subprocess.run(["c++filt", n], capture_output=True, text=True).stdout.strip()  # noqa: E731
WATCH = ["UBLKCP", "UBLKPF", "SYNCS", "UCGABAR", "REDUX", "ATOMS", "ATOMG", "RED", "ATOM", "VOTE", "VOTEU", "MATCH", "SHFL",
         "ACQBULK", "PREEXIT", "LDG", "STG", "LDS", "STS", "LDSM", "MEMBAR", "FENCE", "ERRBAR", "CCTL", "BAR", "POPC", "FLO",
         "DADD", "FADD", "IDP", "LDGSTS", "ELECT", "NANOSLEEP", "LDL", "STL"]
kern, ops, wide = None, {}, {}
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        kern = m.group(1)
        ops[kern], wide[kern] = collections.Counter(), collections.Counter()
        continue
    m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and kern:
        full = m.group(1)
        op = full.split(".")[0]
        ops[kern][op] += 1
        if op in ("LDG", "STG", "LDS", "STS") and (".128" in full or ".64" in full):
            wide[kern][op + (".128" if ".128" in full else ".64")] += 1
print("# static SASS opcode counts per kernel, sm_100a, %s" % os.path.basename(lib))
for k in sorted(ops, key=lambda k: demangle(k)):
    c = ops[k]
    name = demangle(k)
    name = re.sub(r"\(.*", "", name)
    hits = ["%s %d" % (w, c[w]) for w in WATCH if c.get(w)]
    hits += ["%s %d" % (w, n) for w, n in sorted(wide[k].items())]
    print("%-70s total %5d | %s" % (name[:70], sum(c.values()), ", ".join(hits)))
