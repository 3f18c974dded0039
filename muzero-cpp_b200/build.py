"""Build libmzb200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libmzb200.so")
OBJ = os.path.join(HERE, "build")

COMMON = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC",
          "-I", os.path.join(ROOT, "include"), "-I", CSRC]
# (source, extra flags).  tree.cu: no FMA contraction — f64 search arithmetic must round like the x86-64 reference.
SOURCES = [
    ("net_plan.cpp", ["-x", "cu"]),
    ("checkpoint.cpp", ["-x", "cu"]),
    ("net_fp32.cu", []),
    ("net_bf16.cu", []),
    ("tree.cu", ["-fmad=false"]),
    ("engine.cu", []),
    ("engine_search.cu", []),
    ("engine_nccl.cu", []),
    ("value_encoder.cu", ["-fmad=false"]),
    ("selfplay.cu", []),
    ("learn.cu", []),
    ("learn_step.cu", []),
    ("replay.cu", ["-fmad=false"]),
]


def build(verbose=False, force=False):
    os.makedirs(OBJ, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    headers.append(os.path.join(ROOT, "include", "mzb200.h"))
    hdr_time = max(os.path.getmtime(h) for h in headers)
    objs = []
    procs = []
    for src, extra in SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(OBJ, os.path.splitext(src)[0] + ".o")
        objs.append(o)
        if (not force and os.path.exists(o) and os.path.getmtime(o) >= os.path.getmtime(s)
                and os.path.getmtime(o) >= hdr_time):
            continue
        cmd = ["nvcc"] + COMMON + extra + (["-Xptxas", "-v"] if verbose else []) + ["-c", s, "-o", o]
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0 or verbose:
            sys.stderr.write("---- %s\n%s\n" % (src, out))
        failed = failed or p.returncode != 0
    if failed:
        raise RuntimeError("nvcc failed")
    if procs or not os.path.exists(OUT):
        subprocess.check_call(["nvcc", "-shared", "-o", OUT] + objs + ["-lcudart", "-ldl", "-lpthread", "-gencode", "arch=compute_100a,code=sm_100a"])
    return OUT


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv, force="-f" in sys.argv))
