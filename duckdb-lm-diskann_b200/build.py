"""Builds liblmd_b200.so (the C-ABI library: device node store + sm_100a kernels) in-tree with nvcc.

    python duckdb-lm-diskann_b200/build.py [--force]

The .so lands next to this file so that it travels to the GPU box with the repo snapshot.
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "liblmd_b200.so")
SOURCES = ["lmd_store.cu", "lmd_sidecar.cu", "lmd_search_kernels.cu", "lmd_api.cu", "lmd_group.cu", "lmd_rerank_tc.cu", "lmd_build.cu", "lmd_static.cu"]
# lmd_kernel_inst.cu is compiled once per kernel family: (object suffix, extra defines)
INSTANCES = [("trav_m%d_r%d" % (m, r), ["-DLMD_INST_KIND=0", "-DLMD_INST_METRIC=%d" % m, "-DLMD_INST_ROUTE=%d" % r])
             for m, r in ((1, 0), (2, 0), (3, 0), (2, 1), (3, 1), (1, 2), (1, 3), (2, 3), (3, 3))] + \
            [("travf_m%d_r%d" % (m, r), ["-DLMD_INST_KIND=3", "-DLMD_INST_METRIC=%d" % m, "-DLMD_INST_ROUTE=%d" % r])
             for m, r in ((1, 0), (2, 0), (3, 0), (2, 1), (3, 1))] + \
            [("travl_m%d" % m, ["-DLMD_INST_KIND=4", "-DLMD_INST_METRIC=%d" % m]) for m in (1, 2, 3)] + \
            [("travm_m%d" % m, ["-DLMD_INST_KIND=5", "-DLMD_INST_METRIC=%d" % m]) for m in (1, 2, 3)] + \
            [("rerank_m%d" % m, ["-DLMD_INST_KIND=1", "-DLMD_INST_METRIC=%d" % m]) for m in (1, 2, 3)] + \
            [("prune_m%d" % m, ["-DLMD_INST_KIND=2", "-DLMD_INST_METRIC=%d" % m]) for m in (1, 2, 3)]
HEADERS = [os.path.join(CSRC, "lmd_common.cuh"), os.path.join(CSRC, "lmd_kernels.cuh"),
           os.path.join(CSRC, "lmd_build_kernels.cuh"), os.path.join(CSRC, "lmd_plan.hpp"),
           os.path.join(HERE, "..", "include", "lmd_b200.h")]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC,-O3,-Wall,-Wno-unused-function", "-cudart", "static",
]


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    objs, jobs = [], []
    for src in SOURCES:
        sp = os.path.join(CSRC, src)
        op = os.path.join(objdir, src.replace(".cu", ".o"))
        objs.append(op)
        if force or _stale(op, [sp] + HEADERS):
            cmd = [NVCC] + FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", sp, "-o", op]
            jobs.append(cmd)

    sp = os.path.join(CSRC, "lmd_kernel_inst.cu")
    for suffix, defs in INSTANCES:
        op = os.path.join(objdir, "inst_%s.o" % suffix)
        objs.append(op)
        if force or _stale(op, [sp] + HEADERS):
            jobs.append([NVCC] + FLAGS + defs + (["-Xptxas", "-v"] if verbose else []) + ["-c", sp, "-o", op])

    def run(cmd):
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed: %s\n%s\n%s" % (" ".join(cmd), r.stdout, r.stderr))
        return r.stderr

    logs = []
    if jobs:
        with ThreadPoolExecutor(max_workers=min(len(jobs), os.cpu_count() or 4)) as ex:
            logs = list(ex.map(run, jobs))
    if jobs or _stale(OUT, objs):
        run([NVCC, "-shared", "-cudart", "static", "-o", OUT] + objs + ["-lpthread", "-ldl", "-lrt"])
    return OUT, logs


if __name__ == "__main__":
    out, logs = build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    for l in logs:
        if l:
            sys.stderr.write(l)
    print(out)
