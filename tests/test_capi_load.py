"""CPU-side checks of the C-ABI library: it builds, loads, and exports every symbol the headers declare.
No compute call is made here (no GPU in this container)."""
import ctypes as C
import os
import re

import pytest

import h2ot_lib

ROOT = h2ot_lib.ROOT


def _declared(header):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(h2ot_[a-z0-9_]+)\s*\(", src)))


def test_library_builds_and_exports_every_declared_symbol():
    P = h2ot_lib.pkg()
    L = P.load()
    declared = _declared("h2ot.h") + _declared("h2ot_gbm.h")
    assert len(declared) > 50
    for name in declared:
        assert hasattr(L, name), "libh2otree_b200.so does not export %s" % name
    # the Python binding lists the same entry points
    assert sorted(P.ENGINE_SYMBOLS) == _declared("h2ot.h")
    assert sorted(P.GBM_SYMBOLS) == _declared("h2ot_gbm.h")


def test_struct_layouts_match_header_sizes():
    P = h2ot_lib.pkg()
    assert C.sizeof(P.HistDesc) == 40
    assert C.sizeof(P.DecidedNode) == 48
    assert C.sizeof(P.Split) == 32 + 72 + 512
    assert C.sizeof(P.Config) == 64


def test_no_cpu_fallback_without_device():
    P = h2ot_lib.pkg()
    if h2ot_lib.has_gpu():
        pytest.skip("a GPU is present")
    with pytest.raises(P.H2OTError):
        P.Engine(nrows=16, ncols=1)
    assert "no CUDA device" in P.load().h2ot_last_error().decode() or "CUDA" in P.load().h2ot_last_error().decode()


def test_product_does_not_reference_the_oracle():
    pk = os.path.join(ROOT, "h2o-3_b200")
    for dp, _, fs in os.walk(pk):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                txt = open(os.path.join(dp, f), errors="replace").read()
                assert "oracle" not in txt.lower() or f == "__init__.py" and "oracle" not in txt.lower(), f


def test_jni_shim_compiles_against_stub_header():
    """integration/h2otree_jni.c is the C side of the Java binding INTEGRATION.md declares. No JDK exists in this image, so it is
    type-checked against a stand-in jni.h (tests/jni_stub) -- and every native method of the Java class must have its wrapper."""
    import subprocess
    shim = os.path.join(ROOT, "integration", "h2otree_jni.c")
    r = subprocess.run(["gcc", "-std=c11", "-Wall", "-Werror", "-Wno-unused-parameter", "-fsyntax-only", "-I" + os.path.join(ROOT, "tests", "jni_stub"),
                        "-I" + os.path.join(ROOT, "include"), shim], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    java = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    natives = set(re.findall(r"public static native \w+\s+(\w+)\(", java))
    assert len(natives) >= 30
    src = open(shim).read()
    wrapped = set(re.findall(r"Java_hex_tree_gpu_H2OTreeB200_(\w+)\(", src)) | set(re.findall(r"\bFN\((\w+)\)", src))
    assert natives <= wrapped, sorted(natives - wrapped)
