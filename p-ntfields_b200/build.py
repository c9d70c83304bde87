"""Build libpntf_b200.so (sm_100a) in-tree with nvcc.  Usage: python p-ntfields_b200/build.py [--force]"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "pntf", "libpntf_b200.so")
SOURCES = ["capi.cu", "tc_path.cu", "tc_selftest.cu", "mesh_distance.cu"]
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo",
         "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC", "-Xptxas", "-v"]


def newest_source_mtime():
    m = 0.0
    for root, _, files in os.walk(CSRC):
        for f in files:
            m = max(m, os.path.getmtime(os.path.join(root, f)))
    m = max(m, os.path.getmtime(os.path.join(os.path.dirname(HERE), "include", "pntf_b200.h")))
    return m


def build(force=False, verbose=False, defs=(), out=None, objdir="build"):
    """defs / out: experiment builds (scripts/build_variants.py) with extra -D flags into another file."""
    global_out = out or OUT
    if not force and os.path.exists(global_out) and os.path.getmtime(global_out) >= newest_source_mtime():
        return global_out
    objs = []
    os.makedirs(os.path.join(HERE, objdir), exist_ok=True)
    procs = []
    for src in SOURCES:
        obj = os.path.join(HERE, objdir, src.replace(".cu", ".o"))
        objs.append(obj)
        cmd = [NVCC] + FLAGS + list(defs) + ["-c", os.path.join(CSRC, src), "-o", obj]
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for src, p in procs:
        out, _ = p.communicate()
        if verbose or p.returncode != 0:
            print(out)
        if p.returncode != 0:
            raise RuntimeError("nvcc failed on %s" % src)
    cmd = [NVCC, "-shared", "-arch=sm_100a", "-o", global_out] + objs + ["-lcudart"]
    subprocess.check_call(cmd)
    return global_out


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
