#!/usr/bin/env python
"""bench.py — headline benchmark of the scalix-b200 hot path.

    python bench.py --gpus N --steps K --warmup W [--impl reference]

Workload (BASELINE.json configs[1], the configuration the metric is quoted on):
    sclx::algorithm::reduce_last_dim sum over float32 [64, 2^24] (4 GiB) per GPU.
One "step" = one pass of the hot path over that array.  With N > 1 (torchrun, one rank per
GPU) the array is [64, N*2^24] partitioned along its last dimension exactly like
array::set_primary_devices (weak scaling: per-GPU work fixed); each step ends with the real
exchange step of the path — an NCCL all-gather of the N partial rows (N x 256 bytes) folded
in fixed partition order.

value  = algorithmic bytes of all ranks / device time (CUDA events, max over ranks), inputs
         resident in HBM.
e2e    = the same metric through the host-buffer C-ABI call (sclx_b200_reduce_last_dim_host):
         every step copies the 4 GiB input from pinned host memory and reads the 256-byte
         result back, inside the timed region.
The other BASELINE.json configs (scan, execute_kernel axpy, gather) are measured briefly at
N=1 and reported under "also" with their own roofline fractions.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)

ROWS, COLS = 64, 1 << 24
METRIC = "reduce_last_dim effective bandwidth (algorithmic bytes / time)"
UNIT = "GB/s"


def measured_peak():
    try:
        with open(os.path.join(REPO, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """samples nvidia-smi clocks / throttle reasons while a timed region runs"""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.samples, self.stop_flag, self.thread = index, [], False, None

    def _run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                self.samples.append([c.strip() for c in out.strip().split(",")])
            except Exception:
                pass
            time.sleep(0.1)

    def __enter__(self):
        self.thread = threading.Thread(target=self._run, daemon=True)
        self.thread.start()
        return self

    def __exit__(self, *a):
        self.stop_flag = True
        self.thread.join(timeout=6)

    def summary(self):
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for s in self.samples:
            if len(s) < 6:
                continue
            try:
                sm.append(float(s[0]))
                mx = max(mx, float(s[1]))
            except ValueError:
                continue
            for name, v in zip(names, s[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm)}


def host_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_baseline_hostloop(threads):
    """BASELINE.md §4.2 host loop on a bounded sample of the same workload (oracle/ is the
    checker; this is the one leg of bench.py allowed to execute it)."""
    import oracle

    sample_cols = 1 << 22  # [64, 2^22] fp32 = 1 GiB
    a = oracle.uniform_f32_mt(ROWS * sample_cols, 0x5CA11, threads)
    out = np.empty(ROWS, np.float32)
    best = None
    for _ in range(4):
        t0 = time.perf_counter()
        oracle.lib.hostloop_reduce_last_dim_f32(a.ctypes.data_as(C.c_void_p), ROWS, sample_cols,
                                                out.ctypes.data_as(C.c_void_p), threads)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    gbs = a.nbytes / best / 1e9
    return {"value": round(gbs, 2), "unit": UNIT, "cores": threads, "kind": "port",
            "sample": f"host loop over float32 [64, 2^22] (1 GiB of the 4 GiB workload), {threads} pthreads, best of 4"}


# --------------------------------------------------------------------------------------------
def run_reference(args):
    """Reference arm.  Scalix has NO CPU implementation (every array is cudaMallocManaged), so
    this arm runs the UNMODIFIED reference headers through their own public API: the driver
    drivers/parity_driver.cu compiled against /root/reference (oracle/_ref/parity_driver_ref,
    Unified-Memory build) when it was built, else the multithreaded host-loop port."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    n = args.gpus
    threads = host_threads()
    ref_bin = os.path.join(REPO, "oracle", "_ref", "parity_driver_ref")
    line = None
    if os.path.exists(ref_bin):
        env = dict(os.environ)
        env["CUDA_VISIBLE_DEVICES"] = ",".join(str(i) for i in range(n))
        try:
            out = subprocess.run([ref_bin, "bench", "rld_f32", str(COLS * n), str(args.steps), str(args.warmup)],
                                 capture_output=True, text=True, timeout=900, env=env)
            if out.returncode != 0:
                sys.stderr.write(f"reference UM build exited {out.returncode}: {out.stderr[-500:]}\n")
            for ln in out.stdout.splitlines():
                if ln.startswith("{"):
                    r = json.loads(ln)
                    gbs = ROWS * COLS * n * 4 / (r["ms_per_step"] * 1e-3) / 1e9
                    line = {
                        "metric": METRIC, "value": round(gbs, 2), "unit": UNIT, "n_gpus": n, "steps": args.steps,
                        "warmup": args.warmup, "ms_per_step": round(r["ms_per_step"], 4), "higher_is_better": True,
                        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                        "impl": "reference",
                        "config": {"workload": f"reduce_last_dim sum f32 [64, {n}*2^24], reference Unified-Memory build "
                                               "(sclx::algorithm::reduce_last_dim, arrays resident in managed memory)"},
                        "cpu_baseline": {"value": round(gbs, 2), "unit": UNIT, "cores": threads, "kind": "reference",
                                         "sample": "unmodified reference headers, UM GPU build (the reference has no CPU path)"},
                        "e2e": {"value": round(gbs, 2), "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                    }
        except Exception as exc:  # fall through to the port
            sys.stderr.write(f"reference UM build failed: {exc}\n")
    if line is None:
        cb = cpu_baseline_hostloop(threads)
        ms = (ROWS * COLS * 4 / 1e9) / cb["value"] * 1e3
        line = {
            "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": n, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(ms, 3), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "impl": "reference",
            "config": {"workload": "reduce_last_dim sum f32 [64, 2^24]: host-loop port on a 1 GiB sample "
                                   "(the reference has no CPU path; its UM build was not available)"},
            "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        }
    print(json.dumps(line), flush=True)
    return 0


# --------------------------------------------------------------------------------------------
def fill_pinned_uniform(pinned, seed):
    rng = np.random.default_rng(seed)
    a = pinned.array
    chunk = 1 << 24
    for i in range(0, a.size, chunk):
        rng.random(out=a[i:i + chunk], dtype=np.float32)


def time_events(torch, fn, steps):
    """device time of `steps` back-to-back calls of fn() on the current stream, in ms"""
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(steps):
        fn()
    e.record()
    e.synchronize()
    return s.elapsed_time(e)


def also_configs(sclx, torch, stream, peak):
    """the other BASELINE.json configs at N=1, briefly (device-resident, CUDA events)"""
    out = {}

    def entry(name, nbytes, fn, reps=10, warm=3, note=None):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        ms = time_events(torch, fn, reps) / reps
        gbs = nbytes / (ms * 1e-3) / 1e9
        out[name] = {"GB/s": round(gbs, 1), "ms": round(ms, 4), "frac_of_measured_hbm": round(gbs / peak, 3)}
        if note:
            out[name]["note"] = note

    flush = sclx.DeviceBuffer(256 << 20)

    def with_flush(fn):
        # 2^24-element arrays fit in the 126 MB L2: evict them between iterations
        def g():
            sclx.fill(flush, (256 << 20) // 4, 1, np.uint32, stream=stream)
            fn()
        return g

    # config 1: iota + inclusive_scan over 2^24 int32
    n1 = 1 << 24
    a1, o1 = sclx.DeviceBuffer(n1 * 4), sclx.DeviceBuffer(n1 * 4)
    sclx.iota(a1, n1, sclx.I32, 0, stream=stream)
    scan1 = lambda: sclx.inclusive_scan(a1, o1, n1, sclx.I32, sclx.PLUS, stream=stream)
    for _ in range(3):
        scan1()
    torch.cuda.synchronize()
    # time the scan alone with an L2 flush before each (events around the scan only)
    times = []
    for _ in range(10):
        sclx.fill(flush, (256 << 20) // 4, 1, np.uint32, stream=stream)
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); scan1(); e.record(); e.synchronize()
        times.append(s.elapsed_time(e))
    ms = float(np.median(times))
    out["config1_scan_i32_2^24"] = {"GB/s": round(n1 * 8 / (ms * 1e-3) / 1e9, 1), "ms": round(ms, 4),
                                    "frac_of_measured_hbm": round(n1 * 8 / (ms * 1e-3) / 1e9 / peak, 3),
                                    "note": "128 MiB algorithmic, L2 flushed before each launch; launch-latency bound"}
    del a1, o1

    # config 5 (per-GPU share at 8 GPUs): inclusive_scan over 2^29 int64 with a carry-in
    n5 = 1 << 29
    a5, o5 = sclx.DeviceBuffer(n5 * 8), sclx.DeviceBuffer(n5 * 8)
    tot = sclx.DeviceBuffer(64)
    sclx.iota(a5, n5, sclx.I64, 0, stream=stream)
    entry("config5_scan_i64_2^29_per_gpu_share",
          n5 * 16, lambda: sclx.inclusive_scan(a5, o5, n5, sclx.I64, sclx.PLUS, lower_totals=tot.ptr, n_lower=1,
                                               stream=stream), reps=5)
    entry("config5_reduce_pass_i64_2^29", n5 * 8,
          lambda: sclx.reduce(a5, n5, sclx.I64, sclx.PLUS, 0, tot.ptr, stream=stream), reps=5,
          note="pass A (per-partition total) of the multi-device scan")
    del a5, o5

    # config 3: axpy through the execute_kernel range dispatcher over 2^30 f32
    n3 = 1 << 30
    x3, y3 = sclx.DeviceBuffer(n3 * 4), sclx.DeviceBuffer(n3 * 4)
    sclx.fill(x3, n3, 0.5, np.float32, stream=stream)
    entry("config3_axpy_dispatch_f32_2^30", n3 * 12,
          lambda: sclx.axpy_f32(1.0, x3.ptr, y3.ptr, n3, 1, stream=stream), reps=5)
    entry("fill_f32_2^30", n3 * 4, lambda: sclx.fill(y3, n3, 1.0, np.float32, stream=stream), reps=5)

    # config 4 on one GPU: 2^28 random u64 indices into the 2^30-float array
    n4 = 1 << 28
    idx_host = np.random.default_rng(0x6A7).integers(0, n3, size=1 << 24, dtype=np.uint64)
    i4, o4 = sclx.DeviceBuffer(n4 * 8), sclx.DeviceBuffer(n4 * 4)
    for b in range(n4 // idx_host.size):  # replicate a 2^24-index block with a rotating offset
        blk = (idx_host + np.uint64(b * 7919)) % np.uint64(n3)
        sclx.check(sclx.lib.sclx_b200_memcpy_h2d(i4.ptr + b * blk.nbytes, blk.ctypes.data, blk.nbytes, None))
        sclx.synchronize()
    entry("config4_gather_kernel_2^28_into_2^30", n4 * 16,
          lambda: sclx.gather(i4.ptr, n4, 8, [x3], 0, 4, o4.ptr, stream=stream), reps=3,
          note="16 B/gather algorithmic (idx 8 + out 4 + payload 4); HBM moves a 32 B sector per payload")
    entry("config4_gather_via_range_launch", n4 * 16,
          lambda: sclx.gather_launch_f32(i4.ptr, n4, x3.ptr, o4.ptr, 1, stream=stream), reps=3)
    return out


def run_ours(args):
    import torch
    import torch.distributed as dist

    import scalix_b200 as sclx

    n = args.gpus
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if n > 1 and world != n:
        raise SystemExit(f"--gpus {n} needs torchrun with {n} ranks (WORLD_SIZE={world})")
    torch.cuda.set_device(local_rank)
    sclx.check(sclx.lib.sclx_b200_set_device(local_rank))
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    stream = torch.cuda.current_stream().cuda_stream
    peak, peak_kind = measured_peak()
    nbytes = ROWS * COLS * 4

    # ---- synthetic input: this rank's partition, generated in pinned host memory ----------
    pinned = sclx.PinnedBuffer(ROWS * COLS, np.float32)
    fill_pinned_uniform(pinned, 0x5CA11 + rank)
    d_in = sclx.DeviceBuffer(nbytes)
    sclx.check(sclx.lib.sclx_b200_memcpy_h2d(d_in.ptr, pinned.ptr, nbytes, stream))
    row = torch.zeros(ROWS, dtype=torch.float32, device="cuda")
    gathered = torch.zeros(world * ROWS, dtype=torch.float32, device="cuda")
    result = torch.zeros(ROWS, dtype=torch.float32, device="cuda")
    ident = np.zeros(1, np.float32)

    def kernel_only():
        sclx.check(sclx.lib.sclx_b200_reduce_last_dim(d_in.ptr, ROWS, COLS, sclx.F32, sclx.PLUS, ident.ctypes.data,
                                                      row.data_ptr(), stream))

    def exchange():
        if world > 1:
            dist.all_gather_into_tensor(gathered, row)
            sclx.check(sclx.lib.sclx_b200_fold_rows(gathered.data_ptr(), world, ROWS, sclx.F32, sclx.PLUS,
                                                    ident.ctypes.data, result.data_ptr(), stream))

    def step():
        kernel_only()
        exchange()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    launches0 = sclx.launch_count()
    for _ in range(max(args.warmup, 3)):
        step()
    barrier()

    # ---- value: K steps, device time, max over ranks ------------------------------------------
    # the sampler stays on through both timed regions (value and e2e): K device-resident
    # steps last ~12 ms, too short for more than one nvidia-smi sample on their own
    clocks = ClockSampler(local_rank)
    clocks.__enter__()
    barrier()
    l_before = sclx.launch_count()
    total_ms = time_events(torch, step, args.steps)
    barrier()
    gpu_launches = sclx.launch_count() - l_before
    # dominant kernel alone: one event pair per launch
    kt = []
    for _ in range(args.steps):
        kt.append(time_events(torch, kernel_only, 1))
    torch.cuda.synchronize()
    t = torch.tensor([total_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    ms_per_step = total_ms / args.steps
    value = world * nbytes / (ms_per_step * 1e-3) / 1e9
    kernel_ms = float(np.mean(kt))
    achieved = (nbytes + ROWS * 4) / (kernel_ms * 1e-3) / 1e9

    # correctness of what was timed: fp64 check of this rank's row against numpy
    local = row.cpu().numpy().astype(np.float64)
    want = pinned.array.reshape(COLS, ROWS).sum(axis=0, dtype=np.float64)
    if not np.all(np.abs(local - want) <= 1e-5 * np.abs(want)):
        raise SystemExit("bench: reduce_last_dim result outside the 1e-5 tolerance")

    # ---- e2e: host buffers through the C ABI, copies inside the timed region --------------------
    host_out = np.zeros(ROWS, np.float32)
    h_row = torch.zeros(ROWS, dtype=torch.float32).pin_memory()

    def e2e_step():
        sclx.reduce_last_dim_host(pinned.array, ROWS, COLS, sclx.F32, sclx.PLUS, 0.0, host_out)
        if world > 1:
            h_row.copy_(torch.from_numpy(host_out))
            row.copy_(h_row, non_blocking=True)
            exchange()
            h_row.copy_(result)  # blocking D2H of the folded row

    e2e_steps = args.steps
    for _ in range(2):
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    clocks.__exit__()
    t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_s = float(t.item())
    e2e_value = world * nbytes * e2e_steps / e2e_s / 1e9
    if not np.all(np.abs(host_out.astype(np.float64) - want) <= 1e-5 * np.abs(want)) and world == 1:
        raise SystemExit("bench: e2e result outside the 1e-5 tolerance")

    # the bound of the host-buffer path: one plain pinned -> device copy of the same 4 GiB
    def h2d_copy():
        sclx.check(sclx.lib.sclx_b200_memcpy_h2d(d_in.ptr, pinned.ptr, nbytes, stream))

    h2d_copy()
    torch.cuda.synchronize()
    h2d_peak = nbytes / (time_events(torch, h2d_copy, 2) / 2 * 1e-3) / 1e9

    line = None
    if rank == 0:
        line = {
            "metric": METRIC, "value": round(value, 1), "unit": UNIT, "n_gpus": n, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": round(ms_per_step, 4), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {
                "workload": f"reduce_last_dim sum, float32 [64, 2^24] per GPU (global [64, {n}*2^24], last dim "
                            "partitioned like set_primary_devices); N>1 adds an NCCL all-gather of N x 256 B + fixed-order fold",
                "l2": "input (4 GiB per GPU) is larger than the 126 MB L2; no flush needed",
                "peak": f"{peak} GB/s ({peak_kind} HBM copy bandwidth)",
            },
            "roofline": {"bound": "hbm", "achieved": round(achieved, 1), "peak": peak, "unit": "GB/s",
                         "frac": round(achieved / peak, 4), "traffic": None,
                         "kernel": "reduce_last_dim_kernel<float, op_plus, 256, 16, 4, 2>",
                         "kernel_ms": round(kernel_ms, 4)},
            "e2e": {"value": round(e2e_value, 2), "unit": UNIT, "h2d_bytes_per_step": nbytes,
                    "d2h_bytes_per_step": ROWS * 4, "ms_per_step": round(e2e_s / e2e_steps * 1e3, 3),
                    "bound": "pcie h2d", "h2d_copy_gbs_per_gpu": round(h2d_peak, 2),
                    "frac_of_h2d_copy": round(e2e_value / world / h2d_peak, 4)},
            "gpu_launches": int(gpu_launches),
            "clocks": clocks.summary(),
        }
        traffic_file = os.path.join(REPO, "profiles", "traffic.json")
        if os.path.exists(traffic_file):
            try:
                line["roofline"]["traffic"] = json.load(open(traffic_file)).get("reduce_last_dim_bytes_per_launch")
            except Exception:
                pass
    if n == 1:
        pinned.free()
        line["also"] = also_configs(sclx, torch, stream, peak)
        line["cpu_baseline"] = cpu_baseline_hostloop(host_threads())
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank == 0:
        print(json.dumps(line), flush=True)
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
