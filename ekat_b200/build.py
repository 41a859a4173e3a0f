"""Build libekat_b200.so (hand-written CUDA for sm_100a + the C ABI) in-tree with nvcc.

    python -m ekat_b200.build [--force] [--verbose]

nvcc cross-compiles without a GPU, so this runs in the CPU-only build container; the
resulting ekat_b200/lib/libekat_b200.so is git-ignored but travels with the repo to the
GPU box. No torch, no JIT cache: plain `nvcc -shared`.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIBDIR, "libekat_b200.so")
SOURCES = ["capi.cu", "tridiag_levmajor.cu", "tridiag_tma.cu", "tridiag_tma_shared.cu", "tridiag_small.cu", "tridiag_bfb_stream.cu", "tridiag_bfb_shared.cu", "tridiag_tile.cu", "tridiag_cm_rhs.cu", "tridiag_cr_tma.cu", "tridiag_cr_warp.cu", "lin_interp.cu", "lin_interp_lm.cu", "fused.cu", "fused_tma.cu", "util.cu", "view_reduction.cu", "column_ops.cu", "host_pipeline.cu", "no_alloc.cu"]
HEADERS = [os.path.join(CSRC, "common.cuh"), os.path.join(CSRC, "tridiag_lm_common.cuh"), os.path.join(CSRC, "tridiag_tma.cuh"), os.path.join(CSRC, "thomas_two_sided.cuh"), os.path.join(CSRC, "lin_interp_search.cuh"), os.path.join(HERE, "..", "include", "ekat_b200", "ekat_b200.h")]
NVCC = os.environ.get("EKAT_B200_NVCC", "nvcc")
# sm_100a only (no PTX for other archs, no multi-arch fatbin). -lineinfo for ncu source pages.
# No -use_fast_math / -fmad toggles: exactness is spelled with __d*_rn / __f*_rn intrinsics.
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo", "-ccbin", "g++",
         "-Xcompiler", "-fPIC", "-Xptxas", "-v"] + os.environ.get("EKAT_B200_NVCC_FLAGS", "").split()  # (e.g. -DEKB_TRACE_BFB)


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    os.makedirs(LIBDIR, exist_ok=True)
    objs, jobs = [], []
    for s in SOURCES:
        src = os.path.join(CSRC, s)
        obj = os.path.join(LIBDIR, s.replace(".cu", ".o"))
        objs.append(obj)
        if force or _stale(obj, [src] + HEADERS):
            jobs.append([NVCC] + FLAGS + ["-c", src, "-o", obj])

    def run(cmd):
        r = subprocess.run(cmd, capture_output=True, text=True)
        return cmd, r

    logs = []
    with ThreadPoolExecutor(max_workers=min(8, max(1, len(jobs)))) as ex:
        for cmd, r in ex.map(run, jobs):
            logs.append((cmd, r))
            if r.returncode != 0:
                raise RuntimeError("nvcc failed: " + " ".join(cmd) + "\n" + r.stdout + r.stderr)
            if verbose:
                print(" ".join(cmd))
                print(r.stderr)
    if jobs or force or _stale(LIB, objs):
        cmd = [NVCC, "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-ccbin", "g++"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed: " + " ".join(cmd) + "\n" + r.stdout + r.stderr)
    with open(os.path.join(LIBDIR, "ptxas.log"), "a" if not force else "w") as f:
        for cmd, r in logs:
            f.write("### " + " ".join(cmd) + "\n" + r.stderr + "\n")
    return LIB


if __name__ == "__main__":
    path = build(force="--force" in sys.argv, verbose="--verbose" in sys.argv)
    print(path)
