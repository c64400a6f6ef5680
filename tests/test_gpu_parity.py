"""GPU parity tests: every kernel is driven through the C ABI (include/scalix_b200.h) and
compared with the CPU oracle on the same seeded inputs — bit-exact for integer, byte and
index work; within BASELINE.json's 1e-5 relative tolerance (vs the fp64 host sum) for
floating-point sums.  Full-size cases use closed forms the reference's own examples assert.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL_FLOAT_SUM = 1e-5  # BASELINE.json north_star: "within 1e-5 relative of ... an fp64 host sum"


@pytest.fixture(scope="module", autouse=True)
def _device(sclx):
    assert sclx.lib.sclx_b200_device_count() >= 1, sclx.lib.sclx_b200_last_error()
    sclx.check(sclx.lib.sclx_b200_set_device(0))
    before = sclx.launch_count()
    yield
    assert sclx.launch_count() > before, "no kernel of libscalix_b200.so was launched"


def dev(sclx, a):
    return sclx.DeviceBuffer.from_numpy(a)


# ----------------------------------------------------------------------------- fill / iota
@pytest.mark.parametrize("npdt,value", [(np.uint8, 0xAB), (np.uint16, 0xBEEF), (np.int32, -7),
                                        (np.float32, 1.5), (np.float64, -2.25), (np.int64, 1 << 40)])
@pytest.mark.parametrize("n,offset", [(0, 0), (1, 0), (5, 1), (1000, 3), (4097, 0), ((1 << 20) + 13, 2)])
def test_fill(sclx, oracle, npdt, value, n, offset):
    item = np.dtype(npdt).itemsize
    total = n + offset + 8
    buf = sclx.DeviceBuffer(total * item)  # zero-filled
    sclx.fill(buf, n, value, npdt, offset_bytes=offset * item)
    got = buf.to_numpy(npdt)
    want = np.zeros(total, npdt)
    want[offset:offset + n] = oracle.fill(n, npdt, value)
    assert np.array_equal(got, want)


def test_fill_16_byte_elements(sclx):
    v = np.array([0x0123456789ABCDEF, 0x0FEDCBA987654321], np.uint64)
    buf = sclx.DeviceBuffer(16 * 1000)
    sclx.check(sclx.lib.sclx_b200_fill(buf.ptr, 1000, 16, v.ctypes.data, None))
    got = buf.to_numpy(np.uint64).reshape(1000, 2)
    assert np.all(got == v)


@pytest.mark.parametrize("dtype,start", [(0, 0), (0, 0x7FFF0000), (1, 5), (2, -3), (3, 1 << 40), (4, 0.5), (5, -1.0)])
@pytest.mark.parametrize("n,first,offset", [(1, 0, 0), (1000, 0, 0), (4099, 12345, 1), (1 << 20, 1 << 33, 3)])
def test_iota(sclx, oracle, dtype, start, n, first, offset):
    npdt = sclx.NP_DTYPES[dtype]
    item = np.dtype(npdt).itemsize
    buf = sclx.DeviceBuffer((n + offset + 4) * item)
    sclx.iota(buf, n, dtype, start, first_index=first, offset_bytes=offset * item)
    got = buf.to_numpy(npdt)[offset:offset + n]
    want = oracle.iota(n, dtype, start, first)
    assert np.array_equal(got.view(np.uint8), want.view(np.uint8))


# ----------------------------------------------------------------------------- inclusive scan
def _scan_inputs(oracle, dtype, n, seed):
    if dtype in (4, 5):
        a = oracle.uniform_f32(n, seed)
        return a if dtype == 4 else a.astype(np.float64)
    raw = oracle.masked_i64(n, seed, 0xFFFF)
    npdt = {0: np.int32, 1: np.uint32, 2: np.int64, 3: np.uint64}[dtype]
    return (raw - 0x7FFF).astype(npdt) if dtype in (0, 2) else raw.astype(npdt)


SCAN_SIZES = [1, 2, 31, 32, 33, 1000, 6143, 6144, 6145, 12288, 100_003, (1 << 20) + 7]


@pytest.mark.parametrize("dtype", [0, 1, 2, 3])
@pytest.mark.parametrize("op", [0, 2, 3])
@pytest.mark.parametrize("n", SCAN_SIZES)
def test_scan_integer_bit_exact(sclx, oracle, dtype, op, n):
    a = _scan_inputs(oracle, dtype, n, seed=n * 31 + dtype)
    d_in, d_out = dev(sclx, a), sclx.DeviceBuffer(a.nbytes)
    sclx.inclusive_scan(d_in, d_out, n, dtype, op)
    got = d_out.to_numpy(a.dtype)
    assert np.array_equal(got, oracle.inclusive_scan(a, dtype, op))


def test_scan_multiplies_wraps(sclx, oracle):
    a = (np.arange(5000, dtype=np.uint32) * 2 + 3).astype(np.uint32)
    d_in, d_out = dev(sclx, a), sclx.DeviceBuffer(a.nbytes)
    sclx.inclusive_scan(d_in, d_out, a.size, sclx.U32, sclx.MULTIPLIES)
    assert np.array_equal(d_out.to_numpy(np.uint32), oracle.inclusive_scan(a, oracle.U32, oracle.MULTIPLIES))


def test_scan_descriptor_words_of_different_widths_never_alias(sclx):
    """A scan of 4-byte elements leaves 8-byte {status<<32 | value} descriptor words behind, a
    scan of 8-byte elements 16-byte {status, value} words.  If both shared one buffer, a stale
    32-bit VALUE could be decoded as a valid status of the other layout (e.g. a running max of
    4*e+1 reads as AGGREGATE of epoch e).  Each width has its own buffer and epoch: interleave
    the two in both orders with values that collide with every epoch in reach."""
    n = 60_000  # >= 6 tiles of the streaming kernel for either width
    ones64 = np.ones(n, dtype=np.int64)
    d64, o64 = dev(sclx, ones64), sclx.DeviceBuffer(ones64.nbytes)
    o32 = sclx.DeviceBuffer(n * 4)
    for e in range(1, 40):
        for state in (1, 2):
            v = np.full(n, 4 * e + state, dtype=np.uint32)
            d32 = dev(sclx, v)
            sclx.inclusive_scan(d32, o32, n, sclx.U32, sclx.MAX)
            sclx.inclusive_scan(d64, o64, n, sclx.I64, sclx.PLUS)
            assert np.array_equal(o32.to_numpy(np.uint32), v)
            assert np.array_equal(o64.to_numpy(np.int64), np.arange(1, n + 1, dtype=np.int64)), (e, state)
    # the other direction: 64-bit prefixes whose HIGH word looks like a status of the 32-bit layout
    ones32 = np.ones(n, dtype=np.uint32)
    d32 = dev(sclx, ones32)
    for e in range(1, 40):
        hi = np.int64(((e << 2) | 2) << 32)
        v = np.zeros(n, dtype=np.int64)
        v[0] = hi
        d64b = dev(sclx, v)
        sclx.inclusive_scan(d64b, o64, n, sclx.I64, sclx.PLUS)
        sclx.inclusive_scan(d32, o32, n, sclx.U32, sclx.PLUS)
        assert np.array_equal(o32.to_numpy(np.uint32), np.arange(1, n + 1, dtype=np.uint32)), e
        assert np.array_equal(o64.to_numpy(np.int64), np.full(n, hi, dtype=np.int64))


def test_two_streams_use_separate_workspaces(sclx, oracle):
    """the C ABI takes arbitrary caller streams: scans and reductions queued on two streams at
    once must not share tickets, partials or look-back descriptors"""
    s = [C.c_void_p(), C.c_void_p()]
    for h in s:
        sclx.check(sclx.lib.sclx_b200_stream_create(C.byref(h)))
    n, rows, cols = (1 << 20) + 11, 48, 20_001
    a = [_scan_inputs(oracle, 2, n, seed=91 + k) for k in range(2)]
    m = [oracle.masked_i64(rows * cols, 7 + k, 0xFFFF).astype(np.uint32) for k in range(2)]
    d_a = [dev(sclx, x) for x in a]
    d_m = [dev(sclx, x) for x in m]
    d_scan = [sclx.DeviceBuffer(x.nbytes) for x in a]
    d_row = [sclx.DeviceBuffer(rows * 4) for _ in range(2)]
    d_tot = [sclx.DeviceBuffer(8) for _ in range(2)]
    for _ in range(20):
        for k in range(2):
            sclx.inclusive_scan(d_a[k], d_scan[k], n, sclx.I64, sclx.PLUS, stream=s[k])
            sclx.reduce_last_dim(d_m[k], rows, cols, sclx.U32, sclx.PLUS, 0, d_row[k], stream=s[k])
            sclx.reduce(d_a[k], n, sclx.I64, sclx.PLUS, 0, d_tot[k].ptr, stream=s[k])
    for h in s:
        sclx.check(sclx.lib.sclx_b200_stream_synchronize(h))
    for k in range(2):
        assert np.array_equal(d_scan[k].to_numpy(np.int64), np.cumsum(a[k]))
        assert np.array_equal(d_row[k].to_numpy(np.uint32, count=rows),
                              m[k].reshape(cols, rows).sum(axis=0, dtype=np.uint64).astype(np.uint32))
        assert int(d_tot[k].to_numpy(np.int64, count=1)[0]) == int(a[k].sum())
    for h in s:
        sclx.check(sclx.lib.sclx_b200_stream_destroy(h))


@pytest.mark.parametrize("n", [0])
def test_scan_empty(sclx, n):
    buf = sclx.DeviceBuffer(64)
    sclx.inclusive_scan(buf, buf, 0, sclx.I32, sclx.PLUS)
    sclx.synchronize()


@pytest.mark.parametrize("dtype", [0, 2])
def test_scan_unaligned_and_in_place(sclx, oracle, dtype):
    n = 300_001
    a = _scan_inputs(oracle, dtype, n + 3, seed=99)
    item = a.dtype.itemsize
    buf = dev(sclx, a)
    # slice starting at element 1 (or 3): base pointer not 16-byte aligned; scan in place
    for off in (1, 3):
        buf2 = dev(sclx, a)
        sclx.inclusive_scan(buf2, buf2, n, dtype, sclx.PLUS, in_offset=off * item, out_offset=off * item)
        got = buf2.to_numpy(a.dtype)
        assert np.array_equal(got[off:off + n], oracle.inclusive_scan(a[off:off + n], dtype, oracle.PLUS))
        assert np.array_equal(got[:off], a[:off]) and np.array_equal(got[off + n:], a[off + n:])
    del buf


@pytest.mark.parametrize("dtype", [4, 5])
def test_scan_float_tolerance(sclx, oracle, dtype):
    n = 200_000
    a = _scan_inputs(oracle, dtype, n, seed=5)
    d_in, d_out = dev(sclx, a), sclx.DeviceBuffer(a.nbytes)
    sclx.inclusive_scan(d_in, d_out, n, dtype, sclx.PLUS)
    got = d_out.to_numpy(a.dtype).astype(np.float64)
    want = np.cumsum(a.astype(np.float64))
    assert np.allclose(got, want, rtol=1e-5 if dtype == 4 else 1e-12)


def test_scan_golden_ones_example(sclx):
    # /root/reference examples/inclusive_scan.cu:56-92 at its exact size
    n = 2 * 2 * 1_000_000
    d_in = sclx.DeviceBuffer(n * 4)
    sclx.fill(d_in, n, 1, np.int32)
    d_out = sclx.DeviceBuffer(n * 4)
    sclx.inclusive_scan(d_in, d_out, n, sclx.I32, sclx.PLUS)
    assert np.array_equal(d_out.to_numpy(np.int32), np.arange(1, n + 1, dtype=np.int32))


def test_scan_config1_full_size_iota(sclx):
    # BASELINE.json configs[0]: iota + inclusive_scan over 2^24 int32; closed form i(i+1)/2 mod 2^32
    n = 1 << 24
    d = sclx.DeviceBuffer(n * 4)
    sclx.iota(d, n, sclx.I32, 0)
    out = sclx.DeviceBuffer(n * 4)
    sclx.inclusive_scan(d, out, n, sclx.I32, sclx.PLUS)
    i = np.arange(n, dtype=np.uint64)
    want = ((i * (i + 1)) // 2).astype(np.uint32).view(np.int32)
    assert np.array_equal(out.to_numpy(np.int32), want)
    # wrap-around variant
    sclx.iota(d, n, sclx.I32, 0x7FFF0000)
    sclx.inclusive_scan(d, out, n, sclx.I32, sclx.PLUS)
    want = (((i * (i + 1)) // 2) + (i + 1) * 0x7FFF0000).astype(np.uint32).view(np.int32)
    assert np.array_equal(out.to_numpy(np.int32), want)


@pytest.mark.parametrize("dtype,op", [(0, 0), (2, 0), (2, 2), (1, 3)])
@pytest.mark.parametrize("ndev", [2, 3, 8])
def test_scan_partitioned_with_carry(sclx, oracle, dtype, op, ndev):
    """multi-device algorithm on logical partitions of one GPU: per-partition total (reduce),
    then ONE scan per partition carrying the lower totals — must equal the reference's
    scan + host chain + fix-up (oracle_inclusive_scan_partitioned)."""
    leading, last = 4, 50_001
    a = _scan_inputs(oracle, dtype, leading * last, seed=17)
    item = a.dtype.itemsize
    parts = sclx.partition(last, ndev)
    d_in, d_out = dev(sclx, a), sclx.DeviceBuffer(a.nbytes)
    totals = sclx.DeviceBuffer(8 * len(parts))
    ident = {0: 0, 2: -(1 << 62), 3: (1 << 32) - 1}[op]
    for d, (s, ln) in enumerate(parts):
        sclx.reduce(d_in, leading * ln, dtype, op, ident, totals.ptr + d * item, in_offset=leading * s * item)
    for d, (s, ln) in enumerate(parts):
        sclx.inclusive_scan(d_in, d_out, leading * ln, dtype, op, lower_totals=totals.ptr, n_lower=d,
                            in_offset=leading * s * item, out_offset=leading * s * item)
    got = d_out.to_numpy(a.dtype)
    want = oracle.inclusive_scan_partitioned(a, [leading * ln for _, ln in parts], dtype, op)
    assert np.array_equal(got, want)
    assert np.array_equal(got, oracle.inclusive_scan(a, dtype, op))


def test_scan_total_out(sclx, oracle):
    a = _scan_inputs(oracle, 2, 77_777, seed=3)
    d_in, d_out, tot = dev(sclx, a), sclx.DeviceBuffer(a.nbytes), sclx.DeviceBuffer(16)
    sclx.inclusive_scan(d_in, d_out, a.size, sclx.I64, sclx.PLUS, total_out=tot.ptr)
    assert tot.to_numpy(np.int64)[0] == a.sum()


# ----------------------------------------------------------------------------- reduce
@pytest.mark.parametrize("dtype,op,ident", [(0, 0, 0), (1, 2, 0), (2, 3, (1 << 62)), (3, 0, 0), (0, 2, -(1 << 31))])
@pytest.mark.parametrize("n,off", [(0, 0), (1, 0), (1000, 1), (65_537, 0), ((1 << 21) + 5, 3)])
def test_reduce_integer(sclx, oracle, dtype, op, ident, n, off):
    a = _scan_inputs(oracle, dtype, n + off, seed=n + 1)
    item = a.dtype.itemsize
    d_in, out = dev(sclx, a), sclx.DeviceBuffer(16)
    sclx.reduce(d_in, n, dtype, op, ident, out.ptr, in_offset=off * item)
    got = out.to_numpy(a.dtype)[0]
    assert got == oracle.reduce(a[off:off + n], dtype, op, ident)


def test_reduce_float_sum(sclx, oracle):
    n = (1 << 22) + 11
    a = oracle.uniform_f32(n, 0xF00)
    d_in, out = dev(sclx, a), sclx.DeviceBuffer(16)
    sclx.reduce(d_in, n, sclx.F32, sclx.PLUS, 0.0, out.ptr)
    got = float(out.to_numpy(np.float32)[0])
    want = oracle.lib.oracle_sum_f32_f64acc(a.ctypes.data_as(C.c_void_p), n)
    assert abs(got - want) <= RTOL_FLOAT_SUM * abs(want)
    # bit-stable run to run (fixed fold order)
    sclx.reduce(d_in, n, sclx.F32, sclx.PLUS, 0.0, out.ptr)
    assert float(out.to_numpy(np.float32)[0]) == got


# ----------------------------------------------------------------------------- reduce_last_dim
RLD_SHAPES = [(1, 1), (1, 1000), (3, 1), (4, 17), (27, 5000), (64, 4096), (64, 4099), (65, 333),
              (256, 100), (260, 77), (1000, 50), (1024, 33), (5000, 7), (64, 0)]


@pytest.mark.parametrize("rows,cols", RLD_SHAPES)
@pytest.mark.parametrize("dtype,op,ident", [(1, 2, 0), (0, 0, 0), (2, 3, 1 << 62)])
def test_reduce_last_dim_integer(sclx, oracle, rows, cols, dtype, op, ident):
    a = _scan_inputs(oracle, dtype, rows * cols, seed=rows * 131 + cols)
    d_in, out = dev(sclx, a), sclx.DeviceBuffer(max(rows * a.dtype.itemsize, 16))
    sclx.reduce_last_dim(d_in, rows, cols, dtype, op, ident, out)
    got = out.to_numpy(a.dtype, count=rows)
    assert np.array_equal(got, oracle.reduce_last_dim(a, rows, [cols], dtype, op, ident))


@pytest.mark.parametrize("rows,cols", [(64, 1 << 16), (64, 4099), (27, 10_000), (3, 100_000), (300, 999)])
def test_reduce_last_dim_float_sum(sclx, oracle, rows, cols):
    a = oracle.uniform_f32(rows * cols, 0x5CA11)
    d_in, out = dev(sclx, a), sclx.DeviceBuffer(rows * 4)
    sclx.reduce_last_dim(d_in, rows, cols, sclx.F32, sclx.PLUS, 0.0, out)
    got = out.to_numpy(np.float32, count=rows).astype(np.float64)
    want = oracle.reduce_last_dim_f64acc(a, rows)
    assert np.all(np.abs(got - want) <= RTOL_FLOAT_SUM * np.abs(want))


def test_reduce_last_dim_unaligned_base(sclx, oracle):
    rows, cols = 64, 1000
    a = _scan_inputs(oracle, 1, rows * cols + 1, seed=8)
    d_in, out = dev(sclx, a), sclx.DeviceBuffer(rows * 4)
    sclx.reduce_last_dim(d_in, rows, cols, sclx.U32, sclx.PLUS, 0, out, in_offset=4)
    assert np.array_equal(out.to_numpy(np.uint32, count=rows),
                          oracle.reduce_last_dim(a[1:], rows, [cols], oracle.U32, oracle.PLUS, 0))


def test_reduce_last_dim_golden_example_full_size(sclx):
    # /root/reference examples/reduce_last_dim.cu:55-84 at its exact size:
    # array<uint,4>{3,3,3,5'000'000}, a[idx] = idx[3]; max == 4'999'999 for all 27 outputs,
    # distributed over 1/2/4/8 logical partitions and folded in fixed order
    rows, L = 27, 5_000_000
    d_a = sclx.DeviceBuffer(rows * L * 4)
    col = np.repeat(np.arange(L, dtype=np.uint32), rows)
    sclx.check(sclx.lib.sclx_b200_memcpy_h2d(d_a.ptr, col.ctypes.data, col.nbytes, None))
    sclx.synchronize()
    for ndev in (1, 2, 4, 8):
        parts = sclx.partition(L, ndev)
        partial = sclx.DeviceBuffer(len(parts) * rows * 4)
        for d, (s, ln) in enumerate(parts):
            sclx.reduce_last_dim(d_a, rows, ln, sclx.U32, sclx.MAX, 0, partial,
                                 in_offset=s * rows * 4, out_offset=d * rows * 4)
        out = sclx.DeviceBuffer(rows * 4)
        sclx.fold_rows(partial.ptr, len(parts), rows, sclx.U32, sclx.MAX, 0, out.ptr)
        assert np.array_equal(out.to_numpy(np.uint32, count=rows), np.full(rows, L - 1, np.uint32))


def test_reduce_last_dim_config2_full_size_property(sclx):
    # BASELINE.json configs[1] shape [64, 2^24] fp32: closed-form input (row s holds (s+1)/64,
    # exactly representable; all partial sums exact in fp32) -> out[s] == 2^24 * (s+1)/64
    rows, cols = 64, 1 << 24
    d_a = sclx.DeviceBuffer(rows * cols * 4)
    pattern = ((np.arange(rows, dtype=np.float32) + 1) / 64).astype(np.float32)
    # build one 64 MiB block on the host and replicate it on the device
    block_cols = 1 << 18
    block = np.tile(pattern, block_cols)
    for b in range(cols // block_cols):
        sclx.check(sclx.lib.sclx_b200_memcpy_h2d(d_a.ptr + b * block.nbytes, block.ctypes.data, block.nbytes, None))
    sclx.synchronize()
    out = sclx.DeviceBuffer(rows * 4)
    sclx.reduce_last_dim(d_a, rows, cols, sclx.F32, sclx.PLUS, 0.0, out)
    got = out.to_numpy(np.float32, count=rows)
    assert np.array_equal(got, (pattern.astype(np.float64) * cols).astype(np.float32))


def test_reduce_last_dim_partitioned_float(sclx, oracle):
    rows, cols = 64, 100_000
    a = oracle.uniform_f32(rows * cols, 0x5CA11)
    d_in = dev(sclx, a)
    want = oracle.reduce_last_dim_f64acc(a, rows)
    for ndev in (2, 8):
        parts = sclx.partition(cols, ndev)
        partial = sclx.DeviceBuffer(len(parts) * rows * 4)
        for d, (s, ln) in enumerate(parts):
            sclx.reduce_last_dim(d_in, rows, ln, sclx.F32, sclx.PLUS, 0.0, partial,
                                 in_offset=s * rows * 4, out_offset=d * rows * 4)
        out = sclx.DeviceBuffer(rows * 4)
        sclx.fold_rows(partial.ptr, len(parts), rows, sclx.F32, sclx.PLUS, 0.0, out.ptr)
        got = out.to_numpy(np.float32, count=rows).astype(np.float64)
        assert np.all(np.abs(got - want) <= RTOL_FLOAT_SUM * np.abs(want))


# ----------------------------------------------------------------------------- gather
@pytest.mark.parametrize("nparts", [1, 2, 4, 8])
def test_gather_cross_partition(sclx, oracle, nparts):
    n_src, n_idx = 1 << 20, (1 << 18) + 5
    src = oracle.uniform_f32(n_src, 0xABC)
    idx = oracle.index_u64(n_idx, 0x6A7, n_src)
    per = (n_src + nparts - 1) // nparts
    bufs = [dev(sclx, src[p * per:(p + 1) * per]) for p in range(nparts)]
    d_idx, d_out = dev(sclx, idx), sclx.DeviceBuffer(n_idx * 4)
    sclx.gather(d_idx.ptr, n_idx, 8, bufs, per if nparts > 1 else 0, 4, d_out.ptr)
    got = d_out.to_numpy(np.float32)
    assert np.array_equal(got.view(np.uint32), oracle.gather(src, idx).view(np.uint32))


def test_gather_through_range_launch(sclx, oracle):
    n_src, n_idx = 1 << 20, 300_007
    src = oracle.uniform_f32(n_src, 0xABD)
    idx = oracle.index_u64(n_idx, 0x6A8, n_src)
    d_src, d_idx, d_out = dev(sclx, src), dev(sclx, idx), sclx.DeviceBuffer(n_idx * 4)
    for nparts in (1, 3):
        sclx.fill(d_out, n_idx, 0, np.uint32)
        sclx.gather_launch_f32(d_idx.ptr, n_idx, d_src.ptr, d_out.ptr, nparts)
        got = d_out.to_numpy(np.float32)
        assert np.array_equal(got.view(np.uint32), oracle.gather(src, idx).view(np.uint32))


# ----------------------------------------------------------------------------- execute_kernel dispatcher
@pytest.mark.parametrize("n,nparts", [(1, 1), (255, 1), (1_000_003, 1), (1_000_003, 3), (1 << 22, 8)])
def test_axpy_through_range_dispatcher(sclx, oracle, n, nparts):
    x, y = oracle.symmetric_f32(n, 1), oracle.symmetric_f32(n, 2)
    d_x, d_y = dev(sclx, x), dev(sclx, y)
    sclx.axpy_f32(1.5, d_x.ptr, d_y.ptr, n, nparts)
    got = d_y.to_numpy(np.float32)
    assert np.array_equal(got.view(np.uint32), oracle.axpy_f32(1.5, x, y).view(np.uint32))


@pytest.mark.parametrize("shape,nparts", [((4, 4, 1000), 1), ((3, 5, 777), 4), ((1, 7, 12_345), 2)])
def test_axpy_rank3_md_index(sclx, oracle, shape, nparts):
    n = shape[0] * shape[1] * shape[2]
    x, y = oracle.symmetric_f32(n, 3), oracle.symmetric_f32(n, 4)
    d_x, d_y = dev(sclx, x), dev(sclx, y)
    sclx.axpy_f32_rank3(-0.75, d_x.ptr, d_y.ptr, shape[0], shape[1], shape[2], nparts)
    got = d_y.to_numpy(np.float32)
    assert np.array_equal(got.view(np.uint32), oracle.axpy_f32(-0.75, x, y).view(np.uint32))


# ----------------------------------------------------------------------------- host-buffer entry points
def test_reduce_last_dim_host_chunks(sclx, oracle):
    rows, cols = 64, (1 << 19) + 77  # > 2 chunks of 64 MiB
    a = oracle.uniform_f32(rows * cols, 0x5CA11)
    out = np.zeros(rows, np.float32)
    sclx.reduce_last_dim_host(a, rows, cols, sclx.F32, sclx.PLUS, 0.0, out)
    want = oracle.reduce_last_dim_f64acc(a, rows)
    assert np.all(np.abs(out.astype(np.float64) - want) <= RTOL_FLOAT_SUM * np.abs(want))
    b = (a * 1e6).astype(np.uint32)
    outu = np.zeros(rows, np.uint32)
    sclx.reduce_last_dim_host(b, rows, cols, sclx.U32, sclx.PLUS, 0, outu)
    assert np.array_equal(outu, oracle.reduce_last_dim(b, rows, [cols], oracle.U32, oracle.PLUS, 0))


def test_inclusive_scan_host_chunks(sclx, oracle):
    n = (40 << 20) + 3  # int32: 160 MiB -> 3 chunks
    a = oracle.masked_i64(n, 0x77, 0xFFFF).astype(np.int32)
    out = np.zeros(n, np.int32)
    sclx.inclusive_scan_host(a, out, n, sclx.I32, sclx.PLUS)
    assert np.array_equal(out, np.cumsum(a.astype(np.uint32), dtype=np.uint32).view(np.int32))


# ----------------------------------------------------------------------------- BASELINE.json full sizes
def test_axpy_and_gather_config3_config4_full_size(sclx, oracle):
    """configs[2] and configs[3] at their full sizes, bit-exact against the oracle on the same
    index-keyed inputs: y = a*x + y over 2^30 f32 through the range dispatcher (8 partitions,
    like 8 devices), then out[i] = src[index[i]] for 2^28 random u64 indices into that 2^30
    array through the gather kernel (8 source partitions) and through the range launch."""
    n, n_idx, chunk = 1 << 30, 1 << 28, 1 << 26
    d_x, d_y = sclx.DeviceBuffer(n * 4), sclx.DeviceBuffer(n * 4)
    for c in range(0, n, chunk):  # generate + upload in 256 MiB chunks (bounded host memory)
        x, y = oracle.symmetric_f32(chunk, 1, first=c), oracle.symmetric_f32(chunk, 2, first=c)
        sclx.check(sclx.lib.sclx_b200_memcpy_h2d(d_x.ptr + c * 4, x.ctypes.data, x.nbytes, None))
        sclx.check(sclx.lib.sclx_b200_memcpy_h2d(d_y.ptr + c * 4, y.ctypes.data, y.nbytes, None))
        sclx.synchronize()
    sclx.axpy_f32(1.5, d_x.ptr, d_y.ptr, n, 8)
    for c in range(0, n, chunk):
        x, y = oracle.symmetric_f32(chunk, 1, first=c), oracle.symmetric_f32(chunk, 2, first=c)
        got = d_y.to_numpy(np.float32, count=chunk, offset_bytes=c * 4)
        assert np.array_equal(got.view(np.uint32), oracle.axpy_f32(1.5, x, y).view(np.uint32)), f"axpy chunk at {c}"
    del d_y

    # gather: the source is x (regenerated on the host once: 4 GiB)
    src = np.empty(n, np.float32)
    for c in range(0, n, chunk):
        src[c:c + chunk] = oracle.symmetric_f32(chunk, 1, first=c)
    d_idx, d_out = sclx.DeviceBuffer(n_idx * 8), sclx.DeviceBuffer(n_idx * 4)
    for c in range(0, n_idx, chunk):
        idx = oracle.index_u64(chunk, 0x6A7, n, first=c)
        sclx.check(sclx.lib.sclx_b200_memcpy_h2d(d_idx.ptr + c * 8, idx.ctypes.data, idx.nbytes, None))
        sclx.synchronize()
    per = n // 8
    parts = [d_x.ptr + p * per * 4 for p in range(8)]
    for how in ("kernel", "launch"):
        sclx.fill(d_out, n_idx, 0, np.uint32)
        if how == "kernel":
            sclx.gather(d_idx.ptr, n_idx, 8, parts, per, 4, d_out.ptr)
        else:
            sclx.gather_launch_f32(d_idx.ptr, n_idx, d_x.ptr, d_out.ptr, 8)
        for c in range(0, n_idx, chunk):
            idx = oracle.index_u64(chunk, 0x6A7, n, first=c)
            got = d_out.to_numpy(np.float32, count=chunk, offset_bytes=c * 4)
            assert np.array_equal(got.view(np.uint32), oracle.gather(src, idx).view(np.uint32)), f"{how} chunk at {c}"


def test_scan_config5_full_size_closed_form(sclx):
    """configs[4]: inclusive_scan over 2^32 int64, partitioned like 8 devices (pass A totals +
    carry-seeded single-pass scans), on one GPU (32 GiB in + 32 GiB out).  Input iota(0), so
    out[i] = i(i+1)/2 mod 2^64: checked at the last element, in windows around every partition
    boundary, and through a checksum of the whole output (our reduce kernel, mod 2^64)
    = (n-1)n(n+1)/6.  The reference itself cannot run this size on one device (32-bit thread
    arithmetic, scalix/kernel_info.cuh:86-101)."""
    n, ndev, item = 1 << 32, 8, 8
    d_in, d_out = sclx.DeviceBuffer(n * item), sclx.DeviceBuffer(n * item)
    sclx.iota(d_in, n, sclx.I64, 0)
    parts = sclx.partition(n, ndev)
    totals = sclx.DeviceBuffer(8 * ndev)
    for d, (s, ln) in enumerate(parts):
        sclx.reduce(d_in, ln, sclx.I64, sclx.PLUS, 0, totals.ptr + d * item, in_offset=s * item)
    for d, (s, ln) in enumerate(parts):
        sclx.inclusive_scan(d_in, d_out, ln, sclx.I64, sclx.PLUS, lower_totals=totals.ptr, n_lower=d,
                            in_offset=s * item, out_offset=s * item)
    m64 = (1 << 64) - 1

    def closed(i):
        return (i * (i + 1) // 2) & m64

    win = 4096
    starts = {0, n - win}
    for s, ln in parts[1:]:
        starts.update({s - win, s, s + 12345})
    for s in sorted(starts):
        got = d_out.to_numpy(np.uint64, count=win, offset_bytes=s * item)
        want = np.array([closed(i) for i in range(s, s + win)], dtype=np.uint64)
        assert np.array_equal(got, want), f"window at {s}"
    chk = sclx.DeviceBuffer(16)
    sclx.reduce(d_out, n, sclx.U64, sclx.PLUS, 0, chk.ptr)
    assert int(chk.to_numpy(np.uint64)[0]) == ((n - 1) * n * (n + 1) // 6) & m64
    # the single-partition (pure single-pass) scan of the same input must give the same checksum
    sclx.inclusive_scan(d_in, d_out, n, sclx.I64, sclx.PLUS)
    sclx.reduce(d_out, n, sclx.U64, sclx.PLUS, 0, chk.ptr)
    assert int(chk.to_numpy(np.uint64)[0]) == ((n - 1) * n * (n + 1) // 6) & m64
    assert int(d_out.to_numpy(np.uint64, count=1, offset_bytes=(n - 1) * item)[0]) == closed(n - 1)


# ----------------------------------------------------------------------------- array pool (memory manager)
def test_array_pool_reuses_a_freed_range_zero_filled(sclx):
    """a freed array stays mapped in the pool; the next array with the same size and placement
    reuses it and must come back zero-filled (fresh managed memory is zero in the reference,
    examples/distributed_indexing.cu:79 relies on it)"""
    lib = sclx.lib
    sclx.check(lib.sclx_b200_rt_init())
    nbytes = 8 << 20
    dev0, zero = C.c_int(0), C.c_size_t(0)

    class Raw:
        def __init__(self, ptr):
            self.ptr = ptr

    base = C.c_void_p()
    sclx.check(lib.sclx_b200_array_create(nbytes, 1, C.byref(dev0), C.byref(zero), C.byref(base)))
    sclx.fill(Raw(base.value), nbytes // 4, 0xDEADBEEF, np.uint32)
    sclx.synchronize()
    first = base.value
    sclx.check(lib.sclx_b200_array_free(base))
    assert lib.sclx_b200_rt_pool_bytes() >= nbytes
    hits = lib.sclx_b200_rt_pool_hits()
    sclx.check(lib.sclx_b200_array_create(nbytes, 1, C.byref(dev0), C.byref(zero), C.byref(base)))
    assert lib.sclx_b200_rt_pool_hits() == hits + 1 and base.value == first
    got = np.empty(nbytes // 4, np.uint32)
    sclx.check(lib.sclx_b200_memcpy_d2h(got.ctypes.data, base.value, nbytes, None))
    sclx.synchronize()
    assert not got.any()
    sclx.check(lib.sclx_b200_array_free(base))
    sclx.check(lib.sclx_b200_rt_pool_release())
    assert lib.sclx_b200_rt_pool_bytes() == 0
