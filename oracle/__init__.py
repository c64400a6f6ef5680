"""ctypes loader for the CPU oracle (TEST INFRASTRUCTURE — see oracle/scalix_oracle.c).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
may import this package.  The product package (scalix_b200) never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "liboracle.so")

I32, U32, I64, U64, F32, F64 = range(6)
PLUS, MULTIPLIES, MAX, MIN = range(4)
NP_DTYPES = {I32: np.int32, U32: np.uint32, I64: np.int64, U64: np.uint64, F32: np.float32, F64: np.float64}


def build(force=False):
    src = os.path.join(_HERE, "scalix_oracle.c")
    if force or not os.path.exists(_LIB) or os.path.getmtime(_LIB) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "liboracle.so"], stdout=subprocess.DEVNULL)
    return _LIB


def _load():
    build()
    lib = C.CDLL(_LIB)
    u64, i32, vp, f32 = C.c_uint64, C.c_int, C.c_void_p, C.c_float
    sig = {
        "oracle_splitmix64": (u64, [u64, u64]),
        "oracle_fill_uniform_f32": (None, [vp, u64, u64, u64]),
        "oracle_fill_symmetric_f32": (None, [vp, u64, u64, u64]),
        "oracle_fill_index_u64": (None, [vp, u64, u64, u64, u64]),
        "oracle_fill_masked_i64": (None, [vp, u64, u64, u64, u64]),
        "oracle_fill_uniform_f32_mt": (None, [vp, u64, u64, i32]),
        "oracle_as_linear": (u64, [vp, vp, i32]),
        "oracle_create_from_linear": (None, [u64, vp, i32, vp]),
        "oracle_partition": (i32, [u64, i32, vp, vp]),
        "oracle_fill": (None, [vp, u64, i32, vp]),
        "oracle_iota": (i32, [vp, u64, i32, vp, u64]),
        "oracle_inclusive_scan": (i32, [vp, vp, u64, i32, i32]),
        "oracle_inclusive_scan_partitioned": (i32, [vp, vp, vp, i32, i32, i32]),
        "oracle_reduce_last_dim": (i32, [vp, u64, vp, i32, i32, i32, vp, vp]),
        "oracle_reduce_last_dim_f64acc": (None, [vp, u64, u64, vp]),
        "oracle_reduce": (i32, [vp, u64, i32, i32, vp, vp]),
        "oracle_sum_f32_f64acc": (C.c_double, [vp, u64]),
        "oracle_axpy_f32": (None, [f32, vp, vp, u64]),
        "oracle_gather": (None, [vp, vp, u64, i32, vp]),
        "oracle_index_generator_histogram": (None, [vp, u64, i32, vp, vp, vp, i32, vp]),
        "oracle_distributed_indexing": (C.c_double, [vp, u64, vp, vp]),
        "hostloop_reduce_last_dim_f32": (i32, [vp, u64, u64, vp, i32]),
        "hostloop_inclusive_scan_plus": (i32, [vp, vp, u64, i32, i32]),
        "hostloop_axpy_f32": (i32, [f32, vp, vp, u64, i32]),
        "hostloop_gather_f32": (i32, [vp, vp, u64, vp, i32]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype, fn.argtypes = res, args
    return lib


lib = _load()


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _scalar(dtype, v):
    return np.array([v], dtype=NP_DTYPES[dtype])


def uniform_f32(n, seed, first=0):
    out = np.empty(n, np.float32)
    lib.oracle_fill_uniform_f32(_p(out), first, n, seed)
    return out


def uniform_f32_mt(n, seed, threads):
    out = np.empty(n, np.float32)
    lib.oracle_fill_uniform_f32_mt(_p(out), n, seed, threads)
    return out


def symmetric_f32(n, seed, first=0):
    out = np.empty(n, np.float32)
    lib.oracle_fill_symmetric_f32(_p(out), first, n, seed)
    return out


def index_u64(n, seed, modulus, first=0):
    out = np.empty(n, np.uint64)
    lib.oracle_fill_index_u64(_p(out), first, n, seed, modulus)
    return out


def masked_i64(n, seed, mask, first=0):
    out = np.empty(n, np.int64)
    lib.oracle_fill_masked_i64(_p(out), first, n, seed, mask)
    return out


def partition(last_dim, ndev):
    start = np.zeros(max(ndev, 1), np.uint64)
    length = np.zeros(max(ndev, 1), np.uint64)
    used = lib.oracle_partition(last_dim, ndev, _p(start), _p(length))
    if used < 0:
        raise ValueError("ndev must be >= 1")
    return [(int(start[i]), int(length[i])) for i in range(used)]


def as_linear(idx, shape):
    i = np.asarray(idx, np.uint64)
    s = np.asarray(shape, np.uint64)
    return int(lib.oracle_as_linear(_p(i), _p(s), len(s)))


def create_from_linear(lin, shape):
    s = np.asarray(shape, np.uint64)
    out = np.zeros(len(s), np.uint64)
    lib.oracle_create_from_linear(lin, _p(s), len(s), _p(out))
    return [int(v) for v in out]


def fill(n, dtype_np, value):
    out = np.empty(n, dtype_np)
    v = np.array([value], dtype_np)
    lib.oracle_fill(_p(out), n, out.itemsize, _p(v))
    return out


def iota(n, dtype, start, first=0):
    out = np.empty(n, NP_DTYPES[dtype])
    s = _scalar(dtype, start)
    assert lib.oracle_iota(_p(out), n, dtype, _p(s), first) == 0
    return out


def inclusive_scan(a, dtype, op):
    a = np.ascontiguousarray(a)
    out = np.empty_like(a)
    assert lib.oracle_inclusive_scan(_p(a), _p(out), a.size, dtype, op) == 0
    return out


def inclusive_scan_partitioned(a, part_elems, dtype, op):
    a = np.ascontiguousarray(a)
    out = np.empty_like(a)
    pe = np.asarray(part_elems, np.uint64)
    assert int(pe.sum()) == a.size
    assert lib.oracle_inclusive_scan_partitioned(_p(a), _p(out), _p(pe), len(pe), dtype, op) == 0
    return out


def reduce_last_dim(a, rows, part_cols, dtype, op, identity):
    a = np.ascontiguousarray(a)
    pc = np.asarray(part_cols, np.uint64)
    assert int(pc.sum()) * rows == a.size
    out = np.empty(rows, a.dtype)
    ident = _scalar(dtype, identity)
    assert lib.oracle_reduce_last_dim(_p(a), rows, _p(pc), len(pc), dtype, op, _p(ident), _p(out)) == 0
    return out


def reduce_last_dim_f64acc(a, rows):
    a = np.ascontiguousarray(a, np.float32)
    out = np.empty(rows, np.float64)
    lib.oracle_reduce_last_dim_f64acc(_p(a), rows, a.size // rows, _p(out))
    return out


def reduce(a, dtype, op, identity):
    a = np.ascontiguousarray(a)
    out = np.empty(1, a.dtype)
    ident = _scalar(dtype, identity)
    assert lib.oracle_reduce(_p(a), a.size, dtype, op, _p(ident), _p(out)) == 0
    return out[0]


def axpy_f32(a, x, y):
    y = np.array(y, np.float32, copy=True)
    x = np.ascontiguousarray(x, np.float32)
    lib.oracle_axpy_f32(a, _p(x), _p(y), x.size)
    return y


def gather(src, idx):
    src = np.ascontiguousarray(src)
    idx = np.ascontiguousarray(idx, np.uint64)
    out = np.empty(idx.size, src.dtype)
    lib.oracle_gather(_p(src), _p(idx), idx.size, src.itemsize, _p(out))
    return out


def index_generator_histogram(gen, shape, parts):
    gen = np.ascontiguousarray(gen, np.uint64)
    shape = np.asarray(shape, np.uint64)
    ps = np.asarray([p[0] for p in parts], np.uint64)
    pl = np.asarray([p[1] for p in parts], np.uint64)
    hist = np.zeros(int(np.prod(shape)), np.int32)
    lib.oracle_index_generator_histogram(_p(gen), gen.shape[0], gen.shape[1], _p(shape), _p(ps), _p(pl), len(parts), _p(hist))
    return hist


def distributed_indexing(indices, numelem):
    indices = np.ascontiguousarray(indices, np.uint64)
    arr2 = np.zeros(numelem, np.float32)
    arr3 = np.zeros(numelem, np.float32)
    s = lib.oracle_distributed_indexing(_p(indices), numelem, _p(arr2), _p(arr3))
    return s, arr2, arr3
