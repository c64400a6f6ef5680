#!/usr/bin/env python
"""Latency of the blocking public sclx:: API calls at sizes where nothing but the launcher
matters (thread spawn, stream creation and cudaMemAdvise toggling
per launch in the reference; one persistent stream per device and event ordering here).

    python tools/api_latency.py [--gpus 1] [--out gpurun_out/api_latency.json]

Runs drivers/bin/parity_driver_ours and oracle/_ref/parity_driver_ref (same source) with
`bench <case> <n> <steps> <warmup>`: wall clock (steady_clock) around the blocking call,
microseconds per call, ours and the reference Unified-Memory build side by side.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
from multi_gpu_bench import OURS, REF, run  # noqa: E402

CASES = [
    # name, n, what one step is
    ("axpy_f32", 1024, "execute_kernel(...).get() of a 1024-element lambda launch"),
    ("axpy_f32", 1 << 20, "execute_kernel(...).get() of a 2^20-element lambda launch (12 MiB)"),
    ("fill_iota", 1 << 12, "fill(rank-3 array) + iota(rank-1 array), 4096 elements each (two blocking calls)"),
    ("scan_i32", 1024, "algorithm::inclusive_scan over 1024 int32 into a caller-provided result"),
    ("scan_i32", 1 << 24, "config 1: algorithm::inclusive_scan over 2^24 int32"),
    ("rld_f32", 64, "algorithm::reduce_last_dim over float [64, 64]"),
]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--out", default=None)
    ap.add_argument("--emulate", type=int, default=0,
                    help="ours only: N logical devices on GPU 0 (the host-side cost of N-device launch groups)")
    args = ap.parse_args()
    table = {"n_gpus": args.gpus, "emulated_devices": args.emulate, "steps": args.steps, "warmup": args.warmup,
             "timing": "wall clock around the blocking public API call, microseconds per step", "cases": []}
    for name, n, what in CASES:
        row = {"case": name, "n": n, "step": what}
        for tag, binary in (("ours_us", OURS), ("reference_um_us", REF)):
            if not os.path.exists(binary):
                continue
            if args.emulate and tag != "ours_us":
                continue
            r = run(binary, name, n, args.steps, args.warmup, args.gpus, 300, emulate=args.emulate)
            row[tag] = round(r["ms_per_step"] * 1e3, 2) if "ms_per_step" in r else r
        if isinstance(row.get("ours_us"), float) and isinstance(row.get("reference_um_us"), float):
            row["reference_over_ours"] = round(row["reference_um_us"] / row["ours_us"], 1)
        table["cases"].append(row)
        print(json.dumps(row), flush=True)
    if args.out:
        with open(args.out, "w") as f:
            json.dump(table, f, indent=1)
    return 0


if __name__ == "__main__":
    sys.exit(main())
