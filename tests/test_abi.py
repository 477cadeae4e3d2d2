"""CPU-only checks of the drop-in boundary: the C-ABI library loads and exports every symbol that
include/*.h declares (no compute calls here -- there is no GPU in the build container)."""
import ctypes
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
INCLUDE = os.path.join(ROOT, "include")


def declared_symbols():
    """Function and data names declared in the public headers."""
    funcs, data = set(), set()
    for name in sorted(os.listdir(INCLUDE)):
        if not name.endswith(".h"):
            continue
        text = open(os.path.join(INCLUDE, name)).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        text = re.sub(r"//[^\n]*", "", text)
        for m in re.finditer(r"\b([A-Za-z_][A-Za-z0-9_]*)\s*\(([^;{]*?)\)\s*;", text):
            if m.group(1) not in ("defined", "reduce"):
                funcs.add(m.group(1))
        for m in re.finditer(r"extern\s+const\s+uint\s+([A-Z_]+)\s*;", text):
            data.add(m.group(1))
    return funcs, data


@pytest.fixture(scope="module")
def lib_path():
    from cuda_samples_b200 import build
    path = build.LIB
    if not os.path.exists(path):
        path = build.build()
    return path


def test_library_exports_every_declared_symbol(lib_path):
    funcs, data = declared_symbols()
    assert {"histogram256", "histogram64", "scanExclusiveLarge", "scanExclusiveShort", "mergeSort",
            "b200_sort_kv_async", "b200_reduce_f32_host"} <= funcs
    assert data == {"MAX_BATCH_ELEMENTS", "MIN_SHORT_ARRAY_SIZE", "MAX_SHORT_ARRAY_SIZE", "MIN_LARGE_ARRAY_SIZE",
                    "MAX_LARGE_ARRAY_SIZE"}
    lib = ctypes.CDLL(lib_path)
    missing = [f for f in sorted(funcs | data) if not hasattr(lib, f)]
    assert not missing, missing
    # reduce<T> is C++: the three instantiations the reduction sample links against (reduction_kernel.cu:985-989)
    for mangled in ("_Z6reduceIiEviiiiPT_S1_", "_Z6reduceIfEviiiiPT_S1_", "_Z6reduceIdEviiiiPT_S1_"):
        assert hasattr(lib, mangled), mangled


def test_limit_constants_keep_the_samples_values(lib_path):
    # scan.cu:175-179
    lib = ctypes.CDLL(lib_path)
    want = {"MAX_BATCH_ELEMENTS": 64 * 1048576, "MIN_SHORT_ARRAY_SIZE": 4, "MAX_SHORT_ARRAY_SIZE": 1024,
            "MIN_LARGE_ARRAY_SIZE": 2048, "MAX_LARGE_ARRAY_SIZE": 262144}
    for k, v in want.items():
        assert ctypes.c_uint32.in_dll(lib, k).value == v


def test_python_binding_covers_the_abi(lib_path):
    from cuda_samples_b200 import _lib
    funcs, _ = declared_symbols()
    assert funcs <= set(_lib.SIGNATURES), sorted(funcs - set(_lib.SIGNATURES))
    assert _lib.lib.b200_version() >= 100
    assert _lib.lib.b200_launch_count() == 0


def test_library_is_sm100a_only_and_has_no_cpu_fallback(lib_path):
    out = subprocess.run(["cuobjdump", "-lelf", lib_path], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


def test_dropin_headers_compile_as_c_and_cxx(tmp_path):
    src = tmp_path / "t.c"
    src.write_text('#include "histogram_common.h"\n#include "scan_common.h"\n#include "b200prim.h"\n'
                   "int main(void){return (int)sizeof(uint);}\n")
    subprocess.check_call(["gcc", "-std=c11", "-fsyntax-only", "-I", INCLUDE, str(src)])
    src2 = tmp_path / "t.cpp"
    src2.write_text('#include "histogram_common.h"\n#include "scan_common.h"\n#include "mergeSort_common.h"\n'
                    '#include "reduction.h"\n#include "b200prim.h"\n'
                    "void (*fp)(int,int,int,int,float*,float*) = &reduce<float>;\nint main(){return fp == 0;}\n")
    subprocess.check_call(["g++", "-std=c++17", "-fsyntax-only", "-I", INCLUDE, str(src2)])
