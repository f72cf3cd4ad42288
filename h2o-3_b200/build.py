"""In-tree build of libh2otree_b200.so (sm_100a only; nvcc cross-compiles without a GPU)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
SO = os.path.join(HERE, "libh2otree_b200.so")
SOURCES = ["h2ot_engine.cu", "h2ot_noalloc.cu", "host_gbm.cpp"]
HEADERS = ["common.cuh", "k_level.cuh", "k_sparse.cuh", "k_ingest.cuh", "k_xreduce.cuh", "k_rows.cuh", "k_split.cuh", "../../include/h2ot.h", "../../include/h2ot_gbm.h"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
# -fmad=false: the fp64 arithmetic of DHistogram.bin / findBestSplitPoint must round like the JVM (no FMA contraction)
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-fmad=false", "-Xcompiler",
         "-fPIC,-O2,-ffp-contract=off,-Wall,-Wno-unused-function", "-shared", "-cudart", "static", "-ldl"]


def stale():
    if not os.path.exists(SO):
        return True
    t = os.path.getmtime(SO)
    deps = [os.path.join(CSRC, f) for f in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    if not force and not stale():
        return SO
    cmd = [NVCC] + FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", SO] + [os.path.join(CSRC, f) for f in SOURCES]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("nvcc failed building libh2otree_b200.so")
    if verbose:
        sys.stderr.write(r.stderr)
    return SO


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
