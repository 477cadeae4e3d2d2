#!/usr/bin/env python3
"""Build the checker libraries (TEST INFRASTRUCTURE, never part of the product path).

  oracle/liboracle.so          gcc build of oracle/oracle.c (the CPU restatement).
  oracle/_ref/libref_gold.so   the reference's OWN gold sources, compiled unmodified from where
                               they lie under /root/reference (only possible in the build
                               container; the GPU box uses the prebuilt file, which travels with
                               the gpurun snapshot -- oracle/_ref/ is git-ignored, not
                               gpurun-ignored).  No reference source is copied into the repo.

Flags: -O2 -fwrapv, never -ffast-math (Kahan summation in reduceCPU must survive).
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("CUDA_SAMPLES_REFERENCE", "/root/reference")
REF_SAMPLES = os.path.join(REF, "Samples")

# the reference files that make up the CPU gold of the hot path
REF_GOLD = [
    ("2_Concepts_and_Techniques/histogram", "histogram_gold.cpp"),
    ("2_Concepts_and_Techniques/scan", "scan_gold.cpp"),
    ("0_Introduction/mergeSort", "mergeSort_host.cpp"),
    ("0_Introduction/mergeSort", "mergeSort_validate.cpp"),
    ("2_Concepts_and_Techniques/scalarProd", "scalarProd_cpu.cpp"),  # "next" row: reduction siblings
]


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def build_oracle(force=False):
    src = os.path.join(HERE, "oracle.c")
    out = os.path.join(HERE, "liboracle.so")
    if force or _newer(out, [src]):
        cmd = ["gcc", "-O2", "-fwrapv", "-fno-fast-math", "-ffp-contract=off", "-shared", "-fPIC",
               "-std=c11", "-o", out, src]
        subprocess.check_call(cmd)
    return out


def have_reference():
    return all(os.path.exists(os.path.join(REF_SAMPLES, d, f)) for d, f in REF_GOLD)


def build_ref_gold(force=False):
    """Compile the reference gold files in place -> oracle/_ref/libref_gold.so. Returns path or None."""
    out_dir = os.path.join(HERE, "_ref")
    out = os.path.join(out_dir, "libref_gold.so")
    if not have_reference():
        return out if os.path.exists(out) else None
    srcs = [os.path.join(REF_SAMPLES, d, f) for d, f in REF_GOLD]
    if force or _newer(out, srcs):
        os.makedirs(out_dir, exist_ok=True)
        objs = []
        for (d, f), s in zip(REF_GOLD, srcs):
            o = os.path.join(out_dir, f.replace(".cpp", ".o"))
            subprocess.check_call(["g++", "-O2", "-fwrapv", "-fno-fast-math", "-fPIC", "-c", s,
                                   "-I", os.path.join(REF_SAMPLES, d), "-o", o])
            objs.append(o)
        subprocess.check_call(["g++", "-shared", "-o", out] + objs)
        for o in objs:
            os.remove(o)
    return out


REF_REDUCE_DIR = os.path.join(REF_SAMPLES, "2_Concepts_and_Techniques", "reduction")
CUDA_HOME = os.environ.get("CUDA_HOME", "/usr/local/cuda")


def build_ref_reduce(force=False):
    """reduceCPU<T> is a template inside the sample's driver TU (reduction.cpp:167-180): compile that TU, where
    it lies, through oracle/ref_reduce_wrap.cpp (main renamed, C wrappers) -> oracle/_ref/libref_reduce.so.
    Needs the CUDA headers/runtime of this image (the TU includes helper_cuda.h); cudart is linked statically
    so the file loads anywhere the image does."""
    out_dir = os.path.join(HERE, "_ref")
    out = os.path.join(out_dir, "libref_reduce.so")
    src = os.path.join(REF_REDUCE_DIR, "reduction.cpp")
    wrap = os.path.join(HERE, "ref_reduce_wrap.cpp")
    if not os.path.exists(src):
        return out if os.path.exists(out) else None
    if force or _newer(out, [src, wrap]):
        os.makedirs(out_dir, exist_ok=True)
        subprocess.check_call(["g++", "-O2", "-fwrapv", "-fno-fast-math", "-fPIC", "-shared", "-w", "-o", out, wrap,
                               "-I", REF_REDUCE_DIR, "-I", os.path.join(REF, "Common"),
                               "-I", os.path.join(CUDA_HOME, "include"), "-L", os.path.join(CUDA_HOME, "lib64"),
                               "-lcudart_static", "-ldl", "-lrt", "-lpthread"])
    return out


def main():
    force = "--force" in sys.argv
    print("oracle  :", build_oracle(force))
    print("ref gold:", build_ref_gold(force))
    print("ref reduce:", build_ref_reduce(force))


if __name__ == "__main__":
    main()
