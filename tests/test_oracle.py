"""CPU tests: pin the oracle against every known-answer check the reference's examples
embed and against the committed golden fixtures."""
import json
import os

import numpy as np
import pytest

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def test_index_algebra_roundtrip(oracle):
    # /root/reference scalix/index.cuh:53-64,96-108: index 0 is the fastest
    shape = [3, 4, 5]
    assert oracle.as_linear([1, 0, 0], shape) == 1
    assert oracle.as_linear([0, 1, 0], shape) == 3
    assert oracle.as_linear([0, 0, 1], shape) == 12
    for lin in range(60):
        assert oracle.as_linear(oracle.create_from_linear(lin, shape), shape) == lin
    # the last coordinate absorbs overflow
    assert oracle.create_from_linear(60 + 7, shape) == [1, 2, 5]


def test_partition_matches_reference_split(oracle):
    # /root/reference scalix/array.cuh:471-508: ceil(last/ndev) per device, extras dropped
    assert oracle.partition(10, 4) == [(0, 3), (3, 3), (6, 3), (9, 1)]
    assert oracle.partition(8, 8) == [(i, 1) for i in range(8)]
    assert oracle.partition(3, 8) == [(0, 1), (1, 1), (2, 1)]       # "too many devices"
    assert oracle.partition(9, 4) == [(0, 3), (3, 3), (6, 3)]       # 4th device gets nothing
    assert oracle.partition(1 << 24, 8) == [(i << 21, 1 << 21) for i in range(8)]
    with pytest.raises(ValueError):
        oracle.partition(4, 0)


def test_golden_scan_of_ones(oracle):
    # examples/inclusive_scan.cu:56-92: array<int,3>{2,2,1'000'000} of ones -> result[i] == i+1
    n = 2 * 2 * 1_000_000
    ones = np.ones(n, np.int32)
    want = np.arange(1, n + 1, dtype=np.int32)
    assert np.array_equal(oracle.inclusive_scan(ones, oracle.I32, oracle.PLUS), want)
    # the same through the reference's multi-device algorithm, 1..8 partitions of the last dim
    for ndev in (1, 2, 3, 4, 8):
        elems = [4 * ln for _, ln in oracle.partition(1_000_000, ndev)]
        got = oracle.inclusive_scan_partitioned(ones, elems, oracle.I32, oracle.PLUS)
        assert np.array_equal(got, want), ndev


def test_golden_reduce_last_dim_max(oracle):
    # examples/reduce_last_dim.cu:55-84: a[idx] = idx[3] over {3,3,3,L}; max == L-1 for all 27
    # (L = 5'000'000 in the example; 50'000 here keeps the CPU suite fast, same closed form)
    L, rows = 50_000, 27
    a = np.repeat(np.arange(L, dtype=np.uint32), rows)
    for ndev in (1, 2, 4, 8):
        cols = [ln for _, ln in oracle.partition(L, ndev)]
        got = oracle.reduce_last_dim(a, rows, cols, oracle.U32, oracle.MAX, 0)
        assert np.array_equal(got, np.full(rows, L - 1, np.uint32))


def test_golden_distributed_indexing_sum(oracle):
    # examples/distributed_indexing.cu:146-228: sum(arr3) == 4097 * numelem for ANY index table
    numelem = 2000
    idx = oracle.index_u64(64 * numelem, 0xD157, numelem)
    s, arr2, arr3 = oracle.distributed_indexing(idx, numelem)
    assert s == 4097.0 * numelem
    assert np.all(arr3 == 4097.0)
    assert np.all(arr2 == 64.0 + 64.0 * 4097.0)


def test_golden_index_generator_histogram(oracle):
    # examples/index_generator.cu:72-93: 10'000 generated indices into a {4,4,4} histogram;
    # every index is applied by exactly one partition, so the bins sum to 10'000
    n_gen, shape = 10_000, [4, 4, 4]
    raw = oracle.index_u64(3 * n_gen, 0x1D6, 4).reshape(n_gen, 3)
    for ndev in (1, 2, 4):
        hist = oracle.index_generator_histogram(raw, shape, oracle.partition(4, ndev))
        assert int(hist.sum()) == n_gen
        want = np.bincount(raw[:, 0] + 4 * raw[:, 1] + 16 * raw[:, 2], minlength=64)
        assert np.array_equal(hist, want.astype(np.int32))


def test_scan_wraps_like_the_element_type(oracle):
    # BASELINE.md §4.3 C1 wrap-around variant: iota from 0x7fff0000 overflows int32
    a = oracle.iota(1 << 16, oracle.I32, 0x7FFF0000)
    got = oracle.inclusive_scan(a, oracle.I32, oracle.PLUS)
    want = np.cumsum(a.astype(np.uint32), dtype=np.uint32).view(np.int32)
    assert np.array_equal(got, want)


def test_partitioned_scan_equals_serial_for_associative_ops(oracle):
    rng = np.random.default_rng(7)
    a = rng.integers(-1000, 1000, 10_007, dtype=np.int64)
    for op in (oracle.PLUS, oracle.MAX, oracle.MIN):
        serial = oracle.inclusive_scan(a, oracle.I64, op)
        for ndev in (2, 3, 8):
            elems = [ln for _, ln in oracle.partition(a.size, ndev)]
            assert np.array_equal(oracle.inclusive_scan_partitioned(a, elems, oracle.I64, op), serial)


def test_reduce_last_dim_float_against_f64(oracle):
    rows, cols = 64, 4096
    a = oracle.uniform_f32(rows * cols, 0x5CA11)
    f64 = oracle.reduce_last_dim_f64acc(a, rows)
    ref = a.reshape(cols, rows).astype(np.float64).sum(axis=0)
    assert np.allclose(f64, ref, rtol=1e-12)
    got = oracle.reduce_last_dim(a, rows, [cols], oracle.F32, oracle.PLUS, 0.0)
    assert np.allclose(got, f64, rtol=1e-4)  # serial fp32 accumulation: loose bound only


def test_generators_are_index_keyed(oracle):
    a = oracle.uniform_f32(1000, 0x5CA11)
    b = oracle.uniform_f32(500, 0x5CA11, first=500)
    assert np.array_equal(a[500:], b)
    assert a.min() >= 0.0 and a.max() < 1.0
    s = oracle.symmetric_f32(1000, 3)
    assert s.min() >= -1.0 and s.max() < 1.0


def test_golden_fixtures(oracle):
    """fixtures committed under tests/golden (generated by tests/golden/make_golden.py)"""
    with open(os.path.join(GOLDEN, "golden.json")) as f:
        g = json.load(f)
    assert [oracle.lib.oracle_splitmix64(g["splitmix"]["seed"], i) for i in range(8)] == g["splitmix"]["values"]
    a = oracle.uniform_f32(g["rld"]["rows"] * g["rld"]["cols"], g["rld"]["seed"])
    f64 = oracle.reduce_last_dim_f64acc(a, g["rld"]["rows"])
    assert np.allclose(f64, np.array(g["rld"]["row_sums_f64"]), rtol=1e-13)
    m = oracle.masked_i64(g["scan"]["n"], g["scan"]["seed"], 0xFFFF)
    sc = oracle.inclusive_scan(m, oracle.I64, oracle.PLUS)
    assert int(sc[-1]) == g["scan"]["last"]
    assert [int(v) for v in sc[:: g["scan"]["n"] // 8][:8]] == g["scan"]["samples"]


def test_hostloop_baselines_agree_with_oracle(oracle):
    import ctypes as C

    rows, cols = 64, 2048
    a = oracle.uniform_f32(rows * cols, 1)
    out = np.empty(rows, np.float32)
    oracle.lib.hostloop_reduce_last_dim_f32(a.ctypes.data_as(C.c_void_p), rows, cols,
                                            out.ctypes.data_as(C.c_void_p), 3)
    assert np.allclose(out, oracle.reduce_last_dim_f64acc(a, rows), rtol=1e-5)
    x = np.arange(100_003, dtype=np.uint32)
    y = np.empty_like(x)
    oracle.lib.hostloop_inclusive_scan_plus(x.ctypes.data_as(C.c_void_p), y.ctypes.data_as(C.c_void_p),
                                            x.size, 4, 4)
    assert np.array_equal(y, np.cumsum(x, dtype=np.uint32))


def _reference_goldens():
    import glob
    import os

    here = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref")
    out = []
    for path in sorted(glob.glob(os.path.join(here, "*.bin"))):
        case, n = os.path.basename(path)[:-4].rsplit("_", 1)
        out.append((case, int(n), path))
    return out


@pytest.mark.parametrize("case,n,path", _reference_goldens(), ids=lambda v: str(v) if not str(v).startswith("/") else "")
def test_oracle_matches_outputs_of_the_reference_build(oracle, case, n, path):
    """tests/golden/ref/*.bin are outputs of the UNMODIFIED reference (its Unified-Memory build of
    drivers/parity_driver.cu, run on a B200 by tests/golden/make_reference_golden.sh).  The CPU
    oracle must reproduce them: bit-exact for integer work and data movement, 1e-5 relative
    for the float row sums (BASELINE.json's tolerance)."""
    from test_header_api import CASES, compare, oracle_expected

    _, dtype, tol = CASES[case]
    want = oracle_expected(oracle, case, n)
    assert want is not None, case
    got = np.fromfile(path, dtype=dtype)
    compare(got, want, tol, f"{case} n={n}: reference build output vs CPU oracle")


def test_bench_closed_forms_use_the_oracle_generator(oracle):
    """bench.py checks the driver's results on the N devices it is about to time against closed
    forms it computes itself with numpy (it may not execute oracle/ outside its CPU-baseline
    leg): its generator must be the oracle's, value for value."""
    import importlib.util
    import os

    spec = importlib.util.spec_from_file_location(
        "_bench", os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "bench.py"))
    bench = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(bench)
    n = 10_007
    want = np.array([oracle.lib.oracle_splitmix64(0x5CA5, i) for i in range(n)], dtype=np.uint64)
    assert np.array_equal(bench.splitmix64(0x5CA5, np.arange(n, dtype=np.uint64)), want)
    assert np.array_equal(bench.uniform01(0x5CA11, n).view(np.uint32), oracle.uniform_f32(n, 0x5CA11).view(np.uint32))
    assert np.array_equal((bench.splitmix64(0x6A7, np.arange(n, dtype=np.uint64)) % np.uint64(4 * n)),
                          oracle.index_u64(n, 0x6A7, 4 * n))
