"""sclx_b200_reduce_last_dim_noalloc — the allocation-free variant of reduce_last_dim (scratch in a
file-scope static __device__ array, scalix_b200/csrc/no_alloc.cu; no counterpart in the
reference).

Called through the C ABI and compared with the existing entry point sclx_b200_reduce_last_dim
on the same inputs, two calls in a row (the second call finds the first one's contents in the
shared scratch): bit-exact for integers, within the 1e-5 relative tolerance the float-sum tests
use for floats (the two run the same kernel with the same geometry, so in practice bit-equal).
This file is named so that pytest collects it last; it passes alone in a fresh process and after
the other tests.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

RTOL_FLOAT_SUM = 1e-5  # the tolerance of tests/test_gpu_parity.py


def _identity(np_dtype, value):
    return np.array([value], dtype=np_dtype)


def _both(sclx, a, rows, cols, dtype, op, identity, stream=None):
    """(result of the existing entry point, result of the no-alloc one) for input `a`"""
    npdt = a.dtype
    d_in = sclx.DeviceBuffer.from_numpy(a)
    d_ref, d_new = sclx.DeviceBuffer(rows * npdt.itemsize), sclx.DeviceBuffer(rows * npdt.itemsize)
    ident = _identity(npdt, identity)
    sclx.check(sclx.lib.sclx_b200_reduce_last_dim(d_in.ptr, rows, cols, dtype, op, ident.ctypes.data_as(C.c_void_p),
                                                  d_ref.ptr, stream))
    sclx.check(sclx.lib.sclx_b200_reduce_last_dim_noalloc(d_in.ptr, rows, cols, dtype, op,
                                                          ident.ctypes.data_as(C.c_void_p), d_new.ptr, stream))
    sclx.check(sclx.lib.sclx_b200_stream_synchronize(stream))
    return d_ref.to_numpy(npdt, count=rows), d_new.to_numpy(npdt, count=rows)


@pytest.fixture(scope="module", autouse=True)
def _device(sclx):
    assert sclx.lib.sclx_b200_device_count() >= 1, sclx.lib.sclx_b200_last_error()
    sclx.check(sclx.lib.sclx_b200_set_device(0))


def test_noalloc_matches_reduce_last_dim_two_calls_in_a_row(sclx, oracle):
    # call 1: float32 [64, 40_099] sum (the headline shape in small)
    rows, cols = 64, 40_099
    a = oracle.uniform_f32(rows * cols, 0x5CA11)
    ref, new = _both(sclx, a, rows, cols, sclx.F32, sclx.PLUS, 0.0)
    assert np.all(np.abs(new.astype(np.float64) - ref.astype(np.float64)) <= RTOL_FLOAT_SUM * np.abs(ref))
    want = oracle.reduce_last_dim_f64acc(a, rows)
    assert np.all(np.abs(new.astype(np.float64) - want) <= RTOL_FLOAT_SUM * np.abs(want))
    # call 2, right behind it, another shape, dtype and operator: the scratch still holds call
    # 1's partial rows; integers must be bit-exact
    rows2, cols2 = 27, 50_001
    b = oracle.masked_i64(rows2 * cols2, 0x77, 0xFFFFFF).astype(np.uint32)
    ref2, new2 = _both(sclx, b, rows2, cols2, sclx.U32, sclx.MAX, 0)
    assert np.array_equal(ref2, new2)
    assert np.array_equal(new2, b.reshape(cols2, rows2).max(axis=0))
    # and call 1's problem again: same bits as the first time
    _, again = _both(sclx, a, rows, cols, sclx.F32, sclx.PLUS, 0.0)
    assert np.array_equal(again.view(np.uint32), new.view(np.uint32))


@pytest.mark.parametrize("npdt,dtype,op,identity,rows,cols", [
    (np.int64, 2, 0, 0, 256, 3_001),        # the largest 8-byte row count the scratch admits
    (np.float64, 5, 0, 0.0, 33, 20_000),    # odd row count: scalar path
    (np.int32, 0, 3, 2**31 - 1, 512, 999),  # the largest 4-byte row count; min
    (np.float32, 4, 0, 0.0, 64, 1 << 20),   # 256 MiB: every CTA of the persistent grid takes part
])
def test_noalloc_dtypes_and_limits(sclx, oracle, npdt, dtype, op, identity, rows, cols):
    if np.issubdtype(npdt, np.floating):
        a = oracle.uniform_f32(rows * cols, 0xA11).astype(npdt)
    else:
        a = (oracle.masked_i64(rows * cols, 0xB22, 0xFFFFF) - 0x7FFFF).astype(npdt)
    for _ in range(2):  # two calls in a row
        ref, new = _both(sclx, a, rows, cols, dtype, op, identity)
        if np.issubdtype(npdt, np.floating):
            assert np.all(np.abs(new.astype(np.float64) - ref.astype(np.float64)) <= RTOL_FLOAT_SUM * np.abs(ref))
        else:
            assert np.array_equal(ref, new)


def test_noalloc_on_a_caller_stream(sclx, oracle):
    stream = C.c_void_p()
    sclx.check(sclx.lib.sclx_b200_stream_create(C.byref(stream)))
    rows, cols = 64, 10_000
    a = oracle.uniform_f32(rows * cols, 0x515)
    for _ in range(2):
        ref, new = _both(sclx, a, rows, cols, sclx.F32, sclx.PLUS, 0.0, stream=stream)
        assert np.array_equal(ref.view(np.uint32), new.view(np.uint32))
    sclx.check(sclx.lib.sclx_b200_stream_destroy(stream))


def test_noalloc_rejects_sizes_beyond_the_static_scratch(sclx):
    limit = sclx.lib.sclx_b200_noalloc_max_row_bytes()
    assert limit == 2048
    buf = sclx.DeviceBuffer(4096 * 8)
    ident = _identity(np.float32, 0.0)
    rows = limit // 4 + 1
    rc = sclx.lib.sclx_b200_reduce_last_dim_noalloc(buf.ptr, rows, 2, sclx.F32, sclx.PLUS,
                                                    ident.ctypes.data_as(C.c_void_p), buf.ptr, None)
    assert rc == sclx.E_INVALID_ARGUMENT
    msg = sclx.lib.sclx_b200_last_error().decode()
    assert "2048" in msg and "static scratch" in msg
    ident64 = _identity(np.float64, 0.0)
    rc = sclx.lib.sclx_b200_reduce_last_dim_noalloc(buf.ptr, limit // 8 + 1, 2, sclx.F64, sclx.PLUS,
                                                    ident64.ctypes.data_as(C.c_void_p), buf.ptr, None)
    assert rc == sclx.E_INVALID_ARGUMENT
    # at the limit it runs
    sclx.check(sclx.lib.sclx_b200_reduce_last_dim_noalloc(buf.ptr, limit // 4, 2, sclx.F32, sclx.PLUS,
                                                          ident.ctypes.data_as(C.c_void_p), buf.ptr + 16384, None))
    sclx.synchronize()
