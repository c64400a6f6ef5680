"""Short driver for ncu captures: a few launches of each hot kernel at BASELINE.json sizes.
    python tools/ncu_targets.py [rld] [scan64] [scan32] [axpy] [gather] [fill]
Run plain first, then under `ncu --set full -k regex:<kernel>` (see profiles/README.md)."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import scalix_b200 as sclx

which = set(sys.argv[1:]) or {"rld", "scan64", "scan32", "axpy"}
sclx.check(sclx.lib.sclx_b200_set_device(0))
REPS = 3

if "rld" in which:
    rows, cols = 64, 1 << 24
    a = sclx.DeviceBuffer(rows * cols * 4)
    sclx.fill(a, rows * cols, 0.5, np.float32)
    out = sclx.DeviceBuffer(rows * 4)
    for _ in range(REPS):
        sclx.reduce_last_dim(a, rows, cols, sclx.F32, sclx.PLUS, 0.0, out)
    sclx.synchronize()
    assert np.all(out.to_numpy(np.float32, count=rows) == 0.5 * cols)
    del a
if "scan64" in which:
    n = 1 << 29
    a, o = sclx.DeviceBuffer(n * 8), sclx.DeviceBuffer(n * 8)
    sclx.fill(a, n, 1, np.int64)
    for _ in range(REPS):
        sclx.inclusive_scan(a, o, n, sclx.I64, sclx.PLUS)
    sclx.synchronize()
    assert o.to_numpy(np.int64, count=4, offset_bytes=(n - 4) * 8)[-1] == n
    del a, o
if "scan32" in which:
    n = 1 << 24
    a, o = sclx.DeviceBuffer(n * 4), sclx.DeviceBuffer(n * 4)
    sclx.iota(a, n, sclx.I32, 0)
    for _ in range(REPS):
        sclx.inclusive_scan(a, o, n, sclx.I32, sclx.PLUS)
    sclx.synchronize()
    del a, o
if "axpy" in which:
    n = 1 << 30
    x, y = sclx.DeviceBuffer(n * 4), sclx.DeviceBuffer(n * 4)
    sclx.fill(x, n, 0.5, np.float32)
    for _ in range(REPS):
        sclx.axpy_f32(1.0, x.ptr, y.ptr, n, 1)
    sclx.synchronize()
    del x, y
if "fill" in which:
    n = 1 << 30
    y = sclx.DeviceBuffer(n * 4)
    for _ in range(REPS):
        sclx.fill(y, n, 1.0, np.float32)
    sclx.synchronize()
print("ncu_targets ok", sorted(which), "launches", sclx.launch_count())
