"""Builds libscalix_b200.so (sm_100a) in-tree with nvcc.  Used by __graft_entry__.build()."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
LIB_DIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIB_DIR, "libscalix_b200.so")
OBJ_DIR = os.path.join(REPO, "build", "obj")
SOURCES = ["scalix_b200.cu", "launch_instances.cu", "no_alloc.cu", "runtime.cpp"]
NVCC_FLAGS = [
    "-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
    "--extended-lambda", "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC",
    "-I", os.path.join(REPO, "include"),
]


def _newest_header():
    newest = 0.0
    for root, _, files in os.walk(os.path.join(REPO, "include")):
        for f in files:
            newest = max(newest, os.path.getmtime(os.path.join(root, f)))
    return newest


def build(force=False, verbose=False):
    os.makedirs(LIB_DIR, exist_ok=True)
    os.makedirs(OBJ_DIR, exist_ok=True)
    hdr = _newest_header()
    objs, procs = [], []
    for src in SOURCES:
        path = os.path.join(HERE, "csrc", src)
        obj = os.path.join(OBJ_DIR, src + ".o")
        objs.append(obj)
        stale = force or not os.path.exists(obj) or os.path.getmtime(obj) < max(os.path.getmtime(path), hdr)
        if stale:
            cmd = ["nvcc", *NVCC_FLAGS, "-c", path, "-o", obj]
            if verbose:
                print(" ".join(cmd), file=sys.stderr)
            procs.append((src, subprocess.Popen(cmd)))
    for src, p in procs:
        if p.wait() != 0:
            raise RuntimeError(f"nvcc failed on {src}")
    if procs or not os.path.exists(LIB):
        cmd = ["nvcc", "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB, *objs]
        subprocess.check_call(cmd)
        _check_no_exit_stubs()
    return LIB


def _check_no_exit_stubs():
    """nvcc turns a host-side call of a __device__-only function into exit(1) at run time
    instead of a compile error; refuse to ship a library that contains one."""
    try:
        dis = subprocess.run(["objdump", "-d", LIB], capture_output=True, text=True).stdout
    except FileNotFoundError:
        return
    if "<exit@plt>" in dis:
        os.remove(LIB)
        raise RuntimeError("libscalix_b200.so calls exit(): a __device__ function is used from host code")


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
