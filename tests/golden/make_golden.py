"""Generates tests/golden/golden.json.

The reference (header-only CUDA, no CPU path, no test suite) cannot be executed in the
build container, so these fixtures pin the ORACLE's generators and reductions against an
independent numpy evaluation made once here; the reference's own known-answer checks
(closed forms) are asserted directly in tests/test_oracle.py.
Run:  python tests/golden/make_golden.py
"""
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

M64 = (1 << 64) - 1


def splitmix64(seed, index):
    z = (seed + (index + 1) * 0x9E3779B97F4A7C15) & M64
    z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M64
    z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M64
    return z ^ (z >> 31)


def main():
    seed = 0x5CA11
    rows, cols = 64, 512
    vals = np.array([(splitmix64(seed, i) >> 40) for i in range(rows * cols)], dtype=np.float64) / 16777216.0
    vals32 = vals.astype(np.float32)
    row_sums = vals32.astype(np.float64).reshape(cols, rows).sum(axis=0)
    n, sseed = 1 << 14, 0xC5
    masked = [splitmix64(sseed, i) & 0xFFFF for i in range(n)]
    cs = np.cumsum(np.array(masked, dtype=np.int64))
    golden = {
        "splitmix": {"seed": seed, "values": [splitmix64(seed, i) for i in range(8)]},
        "rld": {"seed": seed, "rows": rows, "cols": cols, "row_sums_f64": [float(v) for v in row_sums]},
        "scan": {"seed": sseed, "n": n, "last": int(cs[-1]), "samples": [int(v) for v in cs[:: n // 8][:8]]},
    }
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden.json"), "w") as f:
        json.dump(golden, f, indent=1)


if __name__ == "__main__":
    main()
