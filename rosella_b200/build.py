"""Build librosella_gpu.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "librosella_gpu.so")
HOST_DIR = os.path.join(HERE, "host")
HOST_LIB = os.path.join(HERE, "librosella_host.so")
HOST_SOURCES = ["kmer_table.cpp", "fastx.cpp", "embedder.cpp", "sketcher.cpp", "recover_graph.cpp", "coverage_table.cpp", "capi.cpp"]

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC,-O3,-Wall", "--expt-relaxed-constexpr"]
# per-file extra flags: the f64 re-evaluation must not contract a*b+c into FMA (reference op order)
EXTRA = {"knn_rescore.cu": ["-fmad=false"], "rank_stat.cu": ["-fmad=false"], "cluster_ops.cu": ["-fmad=false"]}
SOURCES = ["rg_api.cu", "knn_transform.cu", "knn_depth2.cu", "knn_depth_sym.cu", "knn_tnf.cu", "knn_tnf_tc.cu", "knn_rescore.cu", "knn_exact.cu", "knn_search.cu",
           "tnf_count.cu", "tnf_count2.cu", "tnf_api.cu", "graph_ops.cu", "graph_static.cu", "sketch.cu", "rg_comm.cu", "knn_dist.cu", "rg_multi.cu", "rank_stat.cu", "cluster_ops.cu", "tnf_host.cpp"]
TOOLS = {"microbench.cu": "microbench", "microbench2.cu": "microbench2"}
HEADERS = ["rg_internal.cuh", "rg_comm.h", "metric_f64.cuh", "pk.cuh", "knn_tnf_common.cuh", "knn_depth_common.cuh", "tnf_tables.h", os.path.join("..", "..", "include", "rosella_gpu.h")]


def _nvcc() -> str:
    for c in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found")


def _host_cxx() -> list[str]:
    # the image exports CXX=/opt/gcc/bin/g++; nvcc is validated against the system g++
    return ["-ccbin", "/usr/bin/g++"] if os.path.exists("/usr/bin/g++") else []


def _stale(target: str, deps: list[str]) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    nvcc = _nvcc()
    hdrs = [os.path.join(CSRC, h) for h in HEADERS] + [os.path.abspath(__file__)]
    objs, procs = [], []
    for src in SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(OBJ, src + ".o")
        objs.append(o)
        if force or _stale(o, [s] + hdrs):
            cmd = [nvcc, *_host_cxx(), *ARCH, *COMMON, *EXTRA.get(src, []), "-c", s, "-o", o]
            if verbose:
                cmd.insert(1, "-Xptxas=-v")
                print(" ".join(cmd), flush=True)
            procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0 or verbose:
            sys.stderr.write(f"--- {src}\n{out}\n")
        failed |= p.returncode != 0
    if failed:
        raise RuntimeError("nvcc failed")
    if force or procs or _stale(LIB, objs):
        # link to a temporary name and rename: a reader (a gpurun snapshot, a running test) never sees a half-written .so
        cmd = [nvcc, *_host_cxx(), *ARCH, "-shared", "-o", LIB + ".tmp", *objs, "-lpthread", "-ldl"]
        subprocess.check_call(cmd)
        os.replace(LIB + ".tmp", LIB)
    # C++ host mirror of the reference interface (host/rosella_host.hpp), on top of the C ABI
    hsrc = [os.path.join(HOST_DIR, f) for f in HOST_SOURCES]
    hdeps = hsrc + [os.path.join(HOST_DIR, "rosella_host.hpp"), os.path.join(HERE, "..", "include", "rosella_gpu.h"), LIB]
    if force or _stale(HOST_LIB, hdeps):
        subprocess.check_call(["/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++", "-O2", "-std=c++17", "-fPIC",
                               "-Wall", "-shared", "-o", HOST_LIB + ".tmp", *hsrc, "-L" + HERE, "-lrosella_gpu", "-lz",
                               "-Wl,-rpath,$ORIGIN"])
        os.replace(HOST_LIB + ".tmp", HOST_LIB)
    # stand-alone micro-benchmarks (bench.py's live FP32 peak; the issue-model probe)
    for src, exe in TOOLS.items():
        s, e = os.path.join(CSRC, src), os.path.join(OBJ, exe)
        if force or _stale(e, [s]):
            subprocess.check_call([nvcc, *_host_cxx(), *ARCH, "-O3", "-std=c++17", "-lineinfo", s, "-o", e])
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
