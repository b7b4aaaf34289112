"""Build libb200knn.so in-tree for sm_100a (nvcc cross-compiles without a GPU).

    python rag-llm-k8s_b200/build.py [--force] [--verbose]

Output: rag-llm-k8s_b200/b200knn/libb200knn.so (git-ignored; travels to the GPU box).
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "b200knn", "libb200knn.so")
OBJ = os.path.join(HERE, "build")
SOURCES = ["index.cu", "multi.cu", "k1_prep.cu", "k2_scan.cu", "k3_tc.cu", "k5_merge.cu", "noalloc.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "-Xcompiler", "-fPIC,-fvisibility=hidden", "--expt-relaxed-constexpr"]


def _deps(src):
    yield os.path.join(CSRC, src)
    for f in os.listdir(CSRC):
        if f.endswith((".h", ".cuh")):
            yield os.path.join(CSRC, f)
    yield os.path.join(HERE, "..", "include", "b200knn.h")


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    objs, procs = [], []
    for src in SOURCES:
        obj = os.path.join(OBJ, src.replace(".cu", ".o"))
        objs.append(obj)
        newest = max(os.path.getmtime(p) for p in _deps(src))
        if not force and os.path.exists(obj) and os.path.getmtime(obj) >= newest:
            continue
        cmd = [NVCC] + FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", os.path.join(CSRC, src), "-o", obj]
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0 or verbose:
            sys.stderr.write("---- %s ----\n%s\n" % (src, out))
        failed |= p.returncode != 0
    if failed:
        raise RuntimeError("nvcc failed")
    if procs or force or not os.path.exists(OUT):
        # cudart linked statically: the .so loads on a box without libcuda (symbol checks on CPU)
        cmd = [NVCC, "-shared", "-o", OUT] + objs + ["-cudart", "static", "-Xlinker", "--exclude-libs,ALL"]
        subprocess.check_call(cmd)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
