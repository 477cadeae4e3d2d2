#!/usr/bin/env python3
"""Build cuda_samples_b200/libb200prim.so from csrc/*.cu with nvcc, sm_100a only.

In-tree on purpose: the .so is git-ignored but travels with the gpurun snapshot.
    python cuda_samples_b200/build.py [--force] [--verbose]
"""
import concurrent.futures
import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libb200prim.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "--use_fast_math=false",
         "-Xcompiler", "-fPIC,-fvisibility=default", "-Xptxas", "-v", "-I", os.path.join(os.path.dirname(HERE), "include")]
FLAGS = [f for f in FLAGS if f != "--use_fast_math=false"]  # never fast-math: Kahan / IEEE division must survive


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _compile(src, obj, verbose):
    r = subprocess.run([NVCC] + FLAGS + ["-c", src, "-o", obj], capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed on %s:\n%s" % (src, r.stderr))
    with open(obj + ".ptxas.txt", "w") as f:  # registers / spills / smem per kernel, kept for the profiles
        f.write(r.stderr)
    if verbose:
        sys.stderr.write(r.stderr)
    return obj


def build(force=False, verbose=False, trace=False):
    global OBJ, LIB, FLAGS
    extra = os.environ.get("B200_EXTRA_FLAGS")
    if extra:  # tuning variants: B200_EXTRA_FLAGS="-DX=1 ..." B200_VARIANT=name -> variants/lib_<name>.so
        name = os.environ.get("B200_VARIANT", "variant")
        OBJ = os.path.join(HERE, "build_" + name)
        os.makedirs(os.path.join(HERE, "variants"), exist_ok=True)
        LIB = os.path.join(HERE, "variants", "lib_%s.so" % name)
        FLAGS = FLAGS + extra.split()
    if trace:  # debug variant with per-tile pipeline timestamps (profiles/scan_trace.py)
        OBJ = os.path.join(HERE, "build_trace")
        LIB = os.path.join(HERE, "libb200prim_trace.so")
        FLAGS = FLAGS + ["-DB200_SCAN_TRACE"]
    srcs = sorted(glob.glob(os.path.join(CSRC, "*.cu")))
    hdrs = glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(CSRC, "*.h")) + \
        glob.glob(os.path.join(os.path.dirname(HERE), "include", "*.h"))
    os.makedirs(OBJ, exist_ok=True)
    jobs = []
    for s in srcs:
        o = os.path.join(OBJ, os.path.basename(s)[:-3] + ".o")
        if force or _stale(o, [s] + hdrs):
            jobs.append((s, o))
    with concurrent.futures.ThreadPoolExecutor(max_workers=8) as ex:
        list(ex.map(lambda a: _compile(a[0], a[1], verbose), jobs))
    objs = [os.path.join(OBJ, os.path.basename(s)[:-3] + ".o") for s in srcs]
    if force or jobs or _stale(LIB, objs):
        subprocess.check_call([NVCC, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-cudart", "static",
                               "-o", LIB] + objs)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv, trace="--trace" in sys.argv))
