#!/usr/bin/env python
"""Tuning sweep: SCLX_B200_WAVES (resident waves per launch) for the framework-owned streaming
kernels, through the C ABI.  One subprocess per setting (the override is read once)."""
import json
import os
import subprocess
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r'''
import sys, json; sys.path.insert(0, %r)
import numpy as np, torch, scalix_b200 as sclx
sclx.check(sclx.lib.sclx_b200_set_device(0))
stream = torch.cuda.current_stream().cuda_stream
def t(fn, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(reps): fn()
    e.record(); e.synchronize()
    return s.elapsed_time(e) / reps
out = {}
n = 1 << 30
a = sclx.DeviceBuffer(n * 4); b = sclx.DeviceBuffer(64)
ms = t(lambda: sclx.fill(a, n, 1.0, np.float32, stream=stream)); out["fill_f32_2^30"] = [round(ms, 4), round(n * 4 / ms / 1e6, 1)]
ms = t(lambda: sclx.iota(a, n // 2, sclx.I64, 0, stream=stream)); out["iota_i64_2^29"] = [round(ms, 4), round(n * 4 / ms / 1e6, 1)]
ms = t(lambda: sclx.reduce(a, n // 2, sclx.I64, sclx.PLUS, 0, b.ptr, stream=stream)); out["reduce_i64_2^29"] = [round(ms, 4), round(n * 4 / ms / 1e6, 1)]
ms = t(lambda: sclx.reduce_last_dim(a, 64, 1 << 24, sclx.F32, sclx.PLUS, 0.0, b, stream=stream)); out["rld_f32_64x2^24"] = [round(ms, 4), round(n * 4 / ms / 1e6, 1)]
ms = t(lambda: sclx.reduce(a, n, sclx.F32, sclx.PLUS, 0, b.ptr, stream=stream)); out["reduce_f32_2^30"] = [round(ms, 4), round(n * 4 / ms / 1e6, 1)]
print(json.dumps(out))
''' % REPO

for waves in sys.argv[1:] or ["1", "2", "4", "8", "16", "64", "256"]:
    env = dict(os.environ, SCLX_B200_WAVES=waves)
    p = subprocess.run([sys.executable, "-c", CHILD], capture_output=True, text=True, env=env, timeout=300)
    line = [ln for ln in p.stdout.splitlines() if ln.startswith("{")]
    print("waves", waves, line[0] if line else p.stderr[-400:], flush=True)
