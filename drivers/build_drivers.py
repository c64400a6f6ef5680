"""Builds the header-API programs against include/ (sm_100a) into drivers/bin/:
   parity_driver_ours  — drivers/parity_driver.cu (public sclx:: API only)
   parity_driver_ours_checked — the same with -DSCLX_GX_CHECKS (bounds checks in the gather kernels)
   ex_<name>           — the reference's five examples, compiled UNMODIFIED from
                         /root/reference/examples (only when that tree is present; nothing is
                         copied into this repository, the binaries are git-ignored)
Called by __graft_entry__.build()."""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
BIN = os.path.join(HERE, "bin")
LIBDIR = os.path.join(REPO, "scalix_b200", "lib")
REF_EXAMPLES = "/root/reference/examples"
EXAMPLES = ["inclusive_scan", "reduce_last_dim", "index_generator", "distributed_indexing", "scaling_performance"]
# examples/inclusive_scan.cu scans with a functor that loops 100,000 times; inlined at every
# fold site of the scan kernels it keeps ptxas -O3 busy for more than ten minutes.  The example
# is a correctness check, so its device code is assembled at -O1 (21 s); the source is unmodified.
EXAMPLE_FLAGS = {"inclusive_scan": ("-Xptxas", "-O1")}
FLAGS = ["-std=c++17", "-O3", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "--extended-lambda",
         "--expt-relaxed-constexpr", "-Xcudafe", "--diag_suppress=177", "-I", os.path.join(REPO, "include"),
         "-L", LIBDIR, "-lscalix_b200", "-Xlinker", "-rpath", "-Xlinker", "$ORIGIN/../../scalix_b200/lib"]


def newest_dependency(src, extra):
    """mtime of the newest repository file `src` includes (nvcc -M; seconds), so touching a header
    only rebuilds the programs that include it"""
    p = subprocess.run(["nvcc", *FLAGS[:FLAGS.index("-L")], *extra, "-M", src], capture_output=True, text=True)
    newest = os.path.getmtime(src)
    if p.returncode != 0:
        return float("inf")
    for tok in p.stdout.replace("\\\n", " ").split():
        if tok.startswith(REPO) and os.path.exists(tok):
            newest = max(newest, os.path.getmtime(tok))
    return newest


def build_one(src, out, extra=()):
    if os.path.exists(out) and os.path.getmtime(out) >= newest_dependency(src, extra):
        return out, 0, ""
    p = subprocess.run(["nvcc", *FLAGS, *extra, "-o", out, src], capture_output=True, text=True)
    return out, p.returncode, p.stdout + p.stderr


def main(force=False):
    os.makedirs(BIN, exist_ok=True)
    # cereal is the reference's vendored third-party dependency (extern/cereal); only the
    # optional serialize case of the driver needs it
    cereal = "/root/reference/extern/cereal/include"
    extra = ("-I", cereal) if os.path.isdir(cereal) else ()
    jobs = [(os.path.join(HERE, "parity_driver.cu"), os.path.join(BIN, "parity_driver_ours"), extra),
            # the same driver with bounds checks compiled into the gather kernels (SCLX_GX_CHECKS)
            (os.path.join(HERE, "parity_driver.cu"), os.path.join(BIN, "parity_driver_ours_checked"),
             (*extra, "-DSCLX_GX_CHECKS"))]
    if os.path.isdir(REF_EXAMPLES):
        for ex in EXAMPLES:
            jobs.append((os.path.join(REF_EXAMPLES, ex + ".cu"), os.path.join(BIN, "ex_" + ex),
                         ("-I", REF_EXAMPLES, *EXAMPLE_FLAGS.get(ex, ()))))
    if force:
        for _, out, _ in jobs:
            if os.path.exists(out):
                os.remove(out)
    failed = []
    with ThreadPoolExecutor(max_workers=6) as pool:
        for out, rc, log in pool.map(lambda j: build_one(*j), jobs):
            if rc != 0:
                failed.append((out, log))
    for out, log in failed:
        sys.stderr.write(f"FAILED {out}\n{log[-4000:]}\n")
    if failed:
        raise SystemExit(1)


if __name__ == "__main__":
    main(force="--force" in sys.argv)
