"""scalix_b200 — Python (ctypes) binding of the C ABI in include/scalix_b200.h.

The product is the C++/CUDA code: the `sclx::` header API in include/scalix/ on top of
libscalix_b200.so.  This module exists so tests/ and bench.py can drive the SAME entry
points a foreign-language binding would use.  There is no CPU fallback and no path into
oracle/: if the shared library is missing, importing fails loudly.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libscalix_b200.so")

I32, U32, I64, U64, F32, F64 = range(6)
PLUS, MULTIPLIES, MAX, MIN = range(4)
NP_DTYPES = {I32: np.int32, U32: np.uint32, I64: np.int64, U64: np.uint64, F32: np.float32, F64: np.float64}
DTYPE_OF_NP = {np.dtype(v): k for k, v in NP_DTYPES.items()}

E_INVALID_ARGUMENT, E_UNSUPPORTED, E_RUNTIME = -1, -2, -3

if not os.path.exists(LIB_PATH):
    raise ImportError(
        f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
        "(scalix_b200 has no CPU fallback)"
    )
lib = C.CDLL(LIB_PATH)

_vp, _sz, _i, _pp = C.c_void_p, C.c_size_t, C.c_int, C.POINTER(C.c_void_p)
# every symbol include/scalix_b200.h declares: name -> (restype, argtypes)
ABI = {
    "sclx_b200_abi_version": (_i, []),
    "sclx_b200_last_error": (C.c_char_p, []),
    "sclx_b200_device_count": (_i, []),
    "sclx_b200_set_device": (_i, [_i]),
    "sclx_b200_dtype_size": (_sz, [_i]),
    "sclx_b200_partition": (_i, [_sz, _i, _vp, _vp]),
    "sclx_b200_malloc": (_i, [_pp, _sz]),
    "sclx_b200_free": (_i, [_vp]),
    "sclx_b200_host_alloc": (_i, [_pp, _sz]),
    "sclx_b200_host_free": (_i, [_vp]),
    "sclx_b200_memcpy_h2d": (_i, [_vp, _vp, _sz, _vp]),
    "sclx_b200_memcpy_d2h": (_i, [_vp, _vp, _sz, _vp]),
    "sclx_b200_memcpy_d2d": (_i, [_vp, _vp, _sz, _vp]),
    "sclx_b200_stream_create": (_i, [_pp]),
    "sclx_b200_stream_destroy": (_i, [_vp]),
    "sclx_b200_stream_synchronize": (_i, [_vp]),
    "sclx_b200_ipc_get_handle": (_i, [_vp, _vp]),
    "sclx_b200_ipc_open_handle": (_i, [_vp, _pp]),
    "sclx_b200_ipc_close": (_i, [_vp]),
    "sclx_b200_fill": (_i, [_vp, _sz, _i, _vp, _vp]),
    "sclx_b200_iota": (_i, [_vp, _sz, _i, _vp, _sz, _vp]),
    "sclx_b200_inclusive_scan": (_i, [_vp, _vp, _sz, _i, _i, _vp, _i, _vp, _vp]),
    "sclx_b200_reduce": (_i, [_vp, _sz, _i, _i, _vp, _vp, _vp]),
    "sclx_b200_reduce_last_dim": (_i, [_vp, _sz, _sz, _i, _i, _vp, _vp, _vp]),
    "sclx_b200_fold_rows": (_i, [_vp, _i, _sz, _i, _i, _vp, _vp, _vp]),
    "sclx_b200_gather": (_i, [_vp, _sz, _i, _vp, _i, _sz, _i, _vp, _vp]),
    "sclx_b200_axpy_f32": (_i, [C.c_float, _vp, _vp, _sz, _i, _vp]),
    "sclx_b200_axpy_f32_rank3": (_i, [C.c_float, _vp, _vp, _sz, _sz, _sz, _i, _vp]),
    "sclx_b200_gather_launch_f32": (_i, [_vp, _sz, _vp, _vp, _i, _vp]),
    "sclx_b200_reduce_last_dim_host": (_i, [_vp, _sz, _sz, _i, _i, _vp, _vp]),
    "sclx_b200_noalloc_max_row_bytes": (_sz, []),
    "sclx_b200_reduce_last_dim_noalloc": (_i, [_vp, _sz, _sz, _i, _i, _vp, _vp, _vp]),
    "sclx_b200_inclusive_scan_host": (_i, [_vp, _vp, _sz, _i, _i]),
    "sclx_b200_launch_count": (C.c_ulonglong, []),
    "sclx_b200_note_launches": (C.c_ulonglong, [C.c_ulonglong]),
    # runtime under the sclx:: header API
    "sclx_b200_rt_init": (_i, []),
    "sclx_b200_rt_set_emulated_devices": (_i, [_i]),
    "sclx_b200_rt_device_count": (_i, []),
    "sclx_b200_rt_physical_device": (_i, [_i]),
    "sclx_b200_rt_stream": (_vp, [_i]),
    "sclx_b200_rt_sm_count": (_i, [_i]),
    "sclx_b200_array_create": (_i, [_sz, _i, _vp, _vp, _pp]),
    "sclx_b200_array_free": (_i, [_vp]),
    "sclx_b200_array_repartition": (_i, [_vp, _i, _vp, _vp]),
    "sclx_b200_array_owner": (_i, [_vp]),
    "sclx_b200_rt_find_captured_arrays": (_i, [_vp, _sz, _i, _pp, _vp]),
    "sclx_b200_rt_before_device_access": (_i, []),
    "sclx_b200_rt_prefetch_to_host": (_i, [_vp, _sz]),
    "sclx_b200_rt_quiesce": (_i, []),
    "sclx_b200_rt_host_faults": (C.c_ulonglong, []),
    "sclx_b200_rt_begin_launch": (_i, [_i, _vp]),
    "sclx_b200_rt_end_launch": (_i, [_i, _vp]),
    "sclx_b200_rt_stream_wait": (_i, [_i, _i]),
    "sclx_b200_rt_record_done": (_i, [_i]),
    "sclx_b200_rt_wait_done": (_i, [_i, _i]),
    "sclx_b200_rt_signal_done": (_i, [_i]),
    "sclx_b200_rt_sync_devices": (_i, [_i, _vp]),
    "sclx_b200_rt_devices_idle": (_i, [_i, _vp]),
    "sclx_b200_rt_host_flag_acquire": (_i, [_pp, _vp]),
    "sclx_b200_rt_host_flag_release": (_i, [_vp]),
    "sclx_b200_rt_host_flag_wait": (_i, [_vp, C.c_uint, _i, _vp]),
    "sclx_b200_rt_workspace_acquire": (_i, [_i, _pp]),
    "sclx_b200_rt_workspace_release": (_i, [_vp, _i]),
    "sclx_b200_rt_workspace_reduce": (_i, [_vp, _sz, _pp, _pp]),
    "sclx_b200_rt_workspace_scan": (_i, [_vp, _sz, _i, _pp, _vp, _pp]),
    "sclx_b200_rt_scratch_acquire": (_i, [_sz, _pp]),
    "sclx_b200_rt_scratch_release": (_i, [_vp, _i]),
    "sclx_b200_rt_exchange_workspace": (_i, [_i, _sz, _pp]),
    "sclx_b200_rt_pool_bytes": (_sz, []),
    "sclx_b200_rt_pool_bytes_on": (_sz, [_i]),
    "sclx_b200_rt_pool_hits": (C.c_ulonglong, []),
    "sclx_b200_rt_pool_release": (_i, []),
}
for _name, (_res, _args) in ABI.items():
    _fn = getattr(lib, _name)  # AttributeError here = header and library disagree
    _fn.restype, _fn.argtypes = _res, _args


class ScalixError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__(f"[{code}] {msg}")
        self.code = code


class InvalidArgument(ScalixError, ValueError):
    """mirrors the std::invalid_argument throws of the reference"""


def check(rc):
    if rc == 0:
        return
    msg = lib.sclx_b200_last_error().decode(errors="replace")
    if rc == E_INVALID_ARGUMENT:
        raise InvalidArgument(rc, msg)
    raise ScalixError(rc, msg)


def _p(a):
    """pointer of a numpy array / int address / None"""
    if a is None:
        return None
    if isinstance(a, int):
        return C.c_void_p(a)
    if isinstance(a, DeviceBuffer):
        return C.c_void_p(a.ptr)
    return a.ctypes.data_as(C.c_void_p)


def partition(last_dim, ndev):
    """(start, len) per device — array::set_primary_devices split arithmetic"""
    n = max(int(ndev), 1)
    start = (C.c_size_t * n)()
    length = (C.c_size_t * n)()
    used = lib.sclx_b200_partition(last_dim, ndev, start, length)
    if used < 0:
        check(used)
    return [(start[i], length[i]) for i in range(used)]


class DeviceBuffer:
    """One partition's explicit device allocation (zero-filled), freed on GC."""

    def __init__(self, nbytes):
        p = C.c_void_p()
        check(lib.sclx_b200_malloc(C.byref(p), nbytes))
        self.ptr = p.value or 0
        self.nbytes = nbytes

    @classmethod
    def from_numpy(cls, a, stream=None):
        a = np.ascontiguousarray(a)
        buf = cls(a.nbytes)
        if a.nbytes:
            check(lib.sclx_b200_memcpy_h2d(buf.ptr, _p(a), a.nbytes, stream))
            check(lib.sclx_b200_stream_synchronize(stream))
        return buf

    def to_numpy(self, dtype, count=None, offset_bytes=0, stream=None):
        dtype = np.dtype(dtype)
        if count is None:
            count = (self.nbytes - offset_bytes) // dtype.itemsize
        out = np.empty(count, dtype)
        if out.nbytes:
            check(lib.sclx_b200_memcpy_d2h(_p(out), self.ptr + offset_bytes, out.nbytes, stream))
            check(lib.sclx_b200_stream_synchronize(stream))
        return out

    def free(self):
        if self.ptr:
            lib.sclx_b200_free(self.ptr)
            self.ptr = 0

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class PinnedBuffer:
    """Pinned host staging memory exposed as a numpy array."""

    def __init__(self, count, dtype):
        dtype = np.dtype(dtype)
        p = C.c_void_p()
        self.nbytes = int(count) * dtype.itemsize
        check(lib.sclx_b200_host_alloc(C.byref(p), self.nbytes))
        self.ptr = p.value
        raw = (C.c_char * self.nbytes).from_address(self.ptr)
        self.array = np.frombuffer(raw, dtype=dtype, count=count)

    def free(self):
        if self.ptr:
            self.array = None
            lib.sclx_b200_host_free(self.ptr)
            self.ptr = 0


def _scalar(dtype, v):
    return np.array([v], dtype=NP_DTYPES[dtype])


def synchronize(stream=None):
    check(lib.sclx_b200_stream_synchronize(stream))


def fill(out, n, value, dtype_np, stream=None, offset_bytes=0):
    v = np.array([value], dtype=dtype_np)
    check(lib.sclx_b200_fill(out.ptr + offset_bytes, n, v.itemsize, _p(v), stream))


def iota(out, n, dtype, start=0, first_index=0, stream=None, offset_bytes=0):
    s = _scalar(dtype, start)
    check(lib.sclx_b200_iota(out.ptr + offset_bytes, n, dtype, _p(s), first_index, stream))


def inclusive_scan(inp, out, n, dtype, op, lower_totals=None, n_lower=0, total_out=None, stream=None,
                   in_offset=0, out_offset=0):
    check(lib.sclx_b200_inclusive_scan(
        inp.ptr + in_offset, out.ptr + out_offset, n, dtype, op,
        _p(lower_totals), n_lower, _p(total_out), stream))


def reduce(inp, n, dtype, op, identity, out, stream=None, in_offset=0):
    ident = _scalar(dtype, identity)
    check(lib.sclx_b200_reduce(inp.ptr + in_offset, n, dtype, op, _p(ident), _p(out), stream))


def reduce_last_dim(inp, rows, cols, dtype, op, identity, out, stream=None, in_offset=0, out_offset=0):
    ident = _scalar(dtype, identity)
    check(lib.sclx_b200_reduce_last_dim(inp.ptr + in_offset, rows, cols, dtype, op, _p(ident),
                                       out.ptr + out_offset, stream))


def fold_rows(parts, nparts, rows, dtype, op, identity, out, stream=None):
    ident = _scalar(dtype, identity)
    check(lib.sclx_b200_fold_rows(_p(parts), nparts, rows, dtype, op, _p(ident), _p(out), stream))


def gather(idx, n, idx_bytes, src_parts, elems_per_part, elem_bytes, out, stream=None):
    ptrs = (C.c_void_p * len(src_parts))(*[p.ptr if isinstance(p, DeviceBuffer) else int(p) for p in src_parts])
    check(lib.sclx_b200_gather(_p(idx), n, idx_bytes, ptrs, len(src_parts), elems_per_part, elem_bytes,
                              _p(out), stream))


def axpy_f32(a, x, y, n, nparts=1, stream=None):
    """y = a*x + y through the sclx::execute_kernel range dispatcher (config 3)"""
    check(lib.sclx_b200_axpy_f32(a, _p(x), _p(y), n, nparts, stream))


def axpy_f32_rank3(a, x, y, d0, d1, last, nparts=1, stream=None):
    check(lib.sclx_b200_axpy_f32_rank3(a, _p(x), _p(y), d0, d1, last, nparts, stream))


def gather_launch_f32(index, n, src, out, nparts=1, stream=None):
    """out[i] = src[index[i]] through the range dispatcher (config 4)"""
    check(lib.sclx_b200_gather_launch_f32(_p(index), n, _p(src), _p(out), nparts, stream))


def reduce_last_dim_host(host_in, rows, cols, dtype, op, identity, host_out):
    ident = _scalar(dtype, identity)
    check(lib.sclx_b200_reduce_last_dim_host(_p(host_in), rows, cols, dtype, op, _p(ident), _p(host_out)))


def inclusive_scan_host(host_in, host_out, n, dtype, op):
    check(lib.sclx_b200_inclusive_scan_host(_p(host_in), _p(host_out), n, dtype, op))


def launch_count():
    return int(lib.sclx_b200_launch_count())
