#!/usr/bin/env python3
"""Comparison harness built from the reference's own sources.  Everything lands in the
git-ignored baseline/_ref/ (travels to the GPU box with the gpurun snapshot); no reference source is copied.

  <sample>_b200   the UNMODIFIED sample driver (main.cpp / reduction.cpp + its CPU gold files) linked against
                  cuda_samples_b200/libb200prim.so instead of the sample's .cu files: the drop-in proof.  The
                  driver generates its own input, runs our kernels, runs its own gold and exits 0 only if they agree.
  <sample>_ref    the sample built entirely from its own sources for sm_100a: "that kernel on the same box".

Only possible where /root/reference exists (the build container).
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
OUT = os.path.join(HERE, "_ref")
REF = os.environ.get("CUDA_SAMPLES_REFERENCE", "/root/reference")
S = os.path.join(REF, "Samples")
CUDA = os.environ.get("CUDA_HOME", "/usr/local/cuda")
LIBDIR = os.path.join(ROOT, "cuda_samples_b200")

SAMPLES = {
    # name: (dir, host sources, device sources)
    "histogram": ("2_Concepts_and_Techniques/histogram", ["main.cpp", "histogram_gold.cpp"],
                  ["histogram64.cu", "histogram256.cu"]),
    "reduction": ("2_Concepts_and_Techniques/reduction", ["reduction.cpp"], ["reduction_kernel.cu"]),
    "scan": ("2_Concepts_and_Techniques/scan", ["main.cpp", "scan_gold.cpp"], ["scan.cu"]),
    "mergeSort": ("0_Introduction/mergeSort", ["main.cpp", "mergeSort_host.cpp", "mergeSort_validate.cpp"],
                  ["mergeSort.cu"]),
    # "next" row
    "sortingNetworks": ("2_Concepts_and_Techniques/sortingNetworks", ["main.cpp", "sortingNetworks_validate.cpp"],
                        ["bitonicSort.cu", "oddEvenMergeSort.cu"]),
    # "next" row rank 4: the kernels live in the driver's own translation unit, so there is no drop-in link;
    # only the reference build, for its own timings on the same box
    "shfl_scan": ("2_Concepts_and_Techniques/shfl_scan", [], ["shfl_scan.cu"]),
    "warpAggregatedAtomicsCG": ("3_CUDA_Features/warpAggregatedAtomicsCG", [], ["warpAggregatedAtomicsCG.cu"]),
}


def have_reference():
    return os.path.isdir(os.path.join(S, "2_Concepts_and_Techniques", "histogram"))


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build_dropin(name):
    d, host, _ = SAMPLES[name]
    src = [os.path.join(S, d, f) for f in host]
    out = os.path.join(OUT, name + "_b200")
    if _stale(out, src + [os.path.join(LIBDIR, "libb200prim.so")]):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-w", "-I", os.path.join(REF, "Common"), "-I", os.path.join(CUDA, "include"),
                               "-I", os.path.join(S, d)] + src +
                              ["-L", LIBDIR, "-lb200prim", "-L", os.path.join(CUDA, "lib64"), "-lcudart",
                               "-Wl,-rpath,$ORIGIN/../../cuda_samples_b200", "-Wl,-rpath," + os.path.join(CUDA, "lib64"),
                               "-o", out])
    return out


def build_reference_gpu(name):
    d, host, dev = SAMPLES[name]
    src = [os.path.join(S, d, f) for f in host + dev]
    out = os.path.join(OUT, name + "_ref")
    if _stale(out, src):
        subprocess.check_call([os.path.join(CUDA, "bin", "nvcc"), "-O3", "-std=c++17", "-w", "-lineinfo",
                               "-gencode", "arch=compute_100a,code=sm_100a", "-I", os.path.join(REF, "Common"),
                               "-I", os.path.join(S, d)] + src + ["-o", out])
    return out


def build_all(which=None):
    if not have_reference():
        return [p for p in (os.path.join(OUT, f) for f in (os.listdir(OUT) if os.path.isdir(OUT) else []))]
    os.makedirs(OUT, exist_ok=True)
    built = []
    for name in SAMPLES:
        if which and name not in which:
            continue
        if SAMPLES[name][1]:
            built.append(build_dropin(name))
        built.append(build_reference_gpu(name))
    return built


if __name__ == "__main__":
    for p in build_all(sys.argv[1:] or None):
        print(p)
