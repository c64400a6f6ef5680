#!/usr/bin/env python
"""In-process multi-GPU measurements through the public sclx:: header API (the reference's own
execution model: ONE process driving N devices), ours and the reference UM build side by side.

    python tools/multi_gpu_bench.py --gpus 2 [--out gpurun_out/multi_n2.json] [--skip-ref]

Runs drivers/bin/parity_driver_ours and oracle/_ref/parity_driver_ref (same source,
drivers/parity_driver.cu) with `bench <case> <n> <steps> <warmup>` on the first N GPUs and
reports wall-clock ms per blocking API call + algorithmic GB/s (SURVEY.md §8d byte counts).
Sizes scale with N the way BASELINE.json's configs do (per-GPU work of configs 2/3/5 fixed).
"""
import argparse
import json
import os
import subprocess
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OURS = os.path.join(REPO, "drivers", "bin", "parity_driver_ours")
REF = os.path.join(REPO, "oracle", "_ref", "parity_driver_ref")


def cases(n_gpus, scale):
    s = scale
    return [
        # name, n argument, algorithmic bytes, note
        ("axpy_f32", (1 << 30) // s, 12 * ((1 << 30) // s), "config 3: y=a*x+y lambda over 2^30 f32 (strong: total fixed)"),
        ("rld_f32", (n_gpus << 24) // s, 4 * 64 * ((n_gpus << 24) // s), "config 2: [64, N*2^24] f32 sum (weak)"),
        ("scan_i64", (n_gpus << 29) // s, 16 * ((n_gpus << 29) // s), "config 5 share: N*2^29 int64 (weak)"),
        ("scan_i32", (1 << 24), 8 * (1 << 24), "config 1: 2^24 int32 (latency bound)"),
        ("gather_f32", (1 << 28) // s, 16 * ((1 << 28) // s), "config 4: 2^28 random indices into 2^30 f32, user lambda (peer loads)"),
        ("gather_x_f32", (1 << 28) // s, 16 * ((1 << 28) // s),
         "config 4 through sclx::algorithm::gather (exchange form when N>1; the reference build runs the lambda)"),
    ]


def run(binary, case, n, steps, warmup, n_gpus, timeout, extra_env=None, emulate=0):
    env = dict(os.environ)
    env.update(extra_env or {})
    env["CUDA_VISIBLE_DEVICES"] = ",".join(str(i) for i in range(n_gpus))
    env.pop("SCLX_B200_EMULATE_DEVICES", None)
    if emulate:  # N logical devices on GPU 0: the host side of an N-device launch on one GPU
        env["SCLX_B200_EMULATE_DEVICES"] = str(emulate)
    try:
        p = subprocess.run([binary, "bench", case, str(n), str(steps), str(warmup)], capture_output=True, text=True,
                           timeout=timeout, env=env)
    except subprocess.TimeoutExpired:
        return {"error": f"timeout after {timeout}s"}
    for ln in p.stdout.splitlines():
        if ln.startswith("{"):
            return json.loads(ln)
    return {"error": f"rc={p.returncode} {p.stderr[-300:]}"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=2)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--scale", type=int, default=1, help="divide every size by this (smoke runs)")
    ap.add_argument("--skip-ref", action="store_true")
    ap.add_argument("--ref-timeout", type=int, default=240)
    ap.add_argument("--out", default=None)
    ap.add_argument("--only", default=None, help="comma-separated case names")
    ap.add_argument("--ref-cases", default=None, help="comma-separated cases the reference build runs (default all)")
    args = ap.parse_args()
    table = {"n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "timing": "wall clock around the blocking public API call",
             "cases": {}}
    for name, n, nbytes, note in cases(args.gpus, args.scale):
        if args.only and name not in args.only.split(","):
            continue
        row = {"n": n, "algorithmic_bytes": nbytes, "note": note}
        if name == "gather_x_f32":
            r = run(OURS, name, n, args.steps, args.warmup, args.gpus, 600, {"SCLX_GATHER_MODE": "direct"})
            if "ms_per_step" in r:
                r["GB/s"] = round(nbytes / (r["ms_per_step"] * 1e-3) / 1e9, 1)
                r["Ggathers/s"] = round(n / (r["ms_per_step"] * 1e-3) / 1e9, 2)
            row["ours_direct_kernel"] = r
            print(name, "ours_direct_kernel", json.dumps(r), flush=True)
        for tag, binary, tmo in (("ours", OURS, 600), ("reference_um", REF, args.ref_timeout)):
            if tag == "reference_um" and (args.skip_ref or not os.path.exists(binary)):
                continue
            if tag == "reference_um" and args.ref_cases and name not in args.ref_cases.split(","):
                continue
            r = run(binary, name, n, args.steps, args.warmup, args.gpus, tmo)
            if "ms_per_step" in r:
                r["GB/s"] = round(nbytes / (r["ms_per_step"] * 1e-3) / 1e9, 1)
                r["GB/s_per_gpu"] = round(r["GB/s"] / args.gpus, 1)
                if name.startswith("gather"):
                    r["Ggathers/s"] = round(n / (r["ms_per_step"] * 1e-3) / 1e9, 2)
            row[tag] = r
            print(name, tag, json.dumps(r), flush=True)
        table["cases"][name] = row
    if args.out:
        with open(args.out, "w") as f:
            json.dump(table, f, indent=1)
    return 0


if __name__ == "__main__":
    sys.exit(main())
