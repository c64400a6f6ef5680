#!/usr/bin/env python
"""Region-size / prefetch sweep of sclx::algorithm::gather (binned form) through the public API.

    python tools/gather_sweep.py [--gpus 1] [--n 268435456] [--out gpurun_out/gather_sweep.json]

One line per setting: wall clock of the blocking call, gathers/s; `direct` = the one-load-per-
index kernel the binned form replaces."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from multi_gpu_bench import OURS, run  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--n", type=int, default=1 << 28)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--regions-kb", default="4096,8192,16384,32768,65536")
    ap.add_argument("--out", default=None)
    args = ap.parse_args()
    rows = []

    def one(tag, env):
        r = run(OURS, "gather_x_f32", args.n, args.steps, 2, args.gpus, 600, env)
        row = {"setting": tag, **r}
        if "ms_per_step" in r:
            row["Ggathers/s"] = round(args.n / (r["ms_per_step"] * 1e-3) / 1e9, 2)
        rows.append(row)
        print(json.dumps(row), flush=True)

    one("direct", {"SCLX_GATHER_MODE": "direct"})
    for kb in args.regions_kb.split(","):
        for pf in (1, 0):
            one(f"binned region={int(kb) >> 10} MiB prefetch={pf}",
                {"SCLX_GATHER_MODE": "exchange", "SCLX_B200_GATHER_REGION_KB": kb, "SCLX_B200_GATHER_PREFETCH": str(pf)})
    if args.out:
        with open(args.out, "w") as f:
            json.dump({"n": args.n, "n_gpus": args.gpus, "rows": rows}, f, indent=1)
    return 0


if __name__ == "__main__":
    sys.exit(main())
