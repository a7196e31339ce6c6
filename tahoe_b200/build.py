"""Build libtahoe_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python -m tahoe_b200.build [--force]
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "lib", "libtahoe_b200.so")
CLI = os.path.join(HERE, "bin", "Tahoe")
SOURCES = ["abi.cu", "packer.cu", "textio.cu", "gen.cu", "kernels_direct.cu",
           "kernels_shared_data.cu", "kernels_smem.cu", "kernels_sparse_smem.cu", "peaks.cu"]
MAX_DEPTH = 14
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC,-O2,-fno-fast-math", "--fmad=true", "-prec-div=true",
              "-prec-sqrt=true", "-ftz=false"]


def _newer(a: str, b: str) -> bool:
    return not os.path.exists(b) or os.path.getmtime(a) > os.path.getmtime(b)


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    os.makedirs(os.path.dirname(LIB), exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    headers.append(os.path.join(HERE, "..", "include", "tahoe_b200.h"))
    hdr_time = max(os.path.getmtime(h) for h in headers)
    jobs = []
    objs = []
    for s in SOURCES:
        src, obj = os.path.join(CSRC, s), os.path.join(OBJ, s + ".o")
        objs.append(obj)
        if force or _newer(src, obj) or hdr_time > os.path.getmtime(obj):
            jobs.append(["nvcc", *NVCC_FLAGS, "-c", src, "-o", obj])
    # k_smem_forest: one object per tree depth, so the instantiations compile in parallel
    src = os.path.join(CSRC, "kernels_smem_depth.cu")
    for d in range(1, MAX_DEPTH + 1):
        obj = os.path.join(OBJ, f"kernels_smem_depth_{d}.o")
        objs.append(obj)
        if force or _newer(src, obj) or hdr_time > os.path.getmtime(obj):
            jobs.insert(0, ["nvcc", *NVCC_FLAGS, f"-DTAHOE_DEPTH={d}", "-c", src, "-o", obj])
    jobs.sort(key=lambda j: 0 if any(a.startswith("-DTAHOE_DEPTH=") for a in j) else 1)

    def run(cmd):
        if verbose:
            print(" ".join(cmd), flush=True)
        subprocess.run(cmd, check=True)

    with ThreadPoolExecutor(max_workers=min(os.cpu_count() or 4, max(1, len(jobs)))) as ex:
        list(ex.map(run, jobs))
    if jobs or not os.path.exists(LIB):
        run(["nvcc", "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB, *objs,
             "-cudart", "static"])
    # the `Tahoe MODEL DATA` command-line clone (include/BaseTahoeTest.h over the C ABI)
    cli_src = os.path.join(HERE, "cli", "main.cpp")
    hdr = os.path.join(HERE, "..", "include", "BaseTahoeTest.h")
    if force or _newer(cli_src, CLI) or _newer(hdr, CLI) or _newer(LIB, CLI):
        os.makedirs(os.path.dirname(CLI), exist_ok=True)
        run(["nvcc", "-O2", "-std=c++17", "-Wno-deprecated-gpu-targets", "-o", CLI, cli_src,
             "-L" + os.path.dirname(LIB), "-ltahoe_b200", "-Xlinker", "-rpath", "-Xlinker", "$ORIGIN/../lib"])
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
