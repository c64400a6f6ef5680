"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol
include/scalix_b200.h declares, and its host-only logic (partition arithmetic, argument
validation) behaves like the reference.  No compute calls: there is no GPU here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(REPO, "include", "scalix_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(sclx_b200_\w+)\s*\(", text)))


def test_every_declared_symbol_is_exported_and_bound(sclx):
    names = _declared_symbols()
    assert len(names) >= 25
    for name in names:
        assert hasattr(sclx.lib, name), f"{name} declared in scalix_b200.h but not exported"
        assert name in sclx.ABI, f"{name} has no ctypes signature in scalix_b200.ABI"
    assert sclx.lib.sclx_b200_abi_version() == 1


def test_python_binding_declares_nothing_extra(sclx):
    assert sorted(sclx.ABI) == _declared_symbols()


def test_partition_matches_oracle(sclx, oracle):
    for last in (0, 1, 3, 9, 10, 64, 1000, 5_000_000, 1 << 24, (1 << 32) + 5):
        for ndev in (1, 2, 3, 4, 7, 8):
            assert sclx.partition(last, ndev) == oracle.partition(last, ndev), (last, ndev)


def test_partition_rejects_zero_devices(sclx):
    with pytest.raises(sclx.InvalidArgument):
        sclx.partition(8, 0)


def test_dtype_sizes(sclx):
    for dt, npdt in sclx.NP_DTYPES.items():
        assert sclx.lib.sclx_b200_dtype_size(dt) == np.dtype(npdt).itemsize
    assert sclx.lib.sclx_b200_dtype_size(99) == 0


def test_argument_errors_do_not_need_a_device(sclx):
    # null pointers / bad dtypes are rejected before any CUDA call, with the error text set
    one = np.zeros(1, np.float32)
    rc = sclx.lib.sclx_b200_inclusive_scan(None, None, 8, sclx.I32, sclx.PLUS, None, 0, None, None)
    assert rc == sclx.E_INVALID_ARGUMENT
    assert b"inclusive_scan" in sclx.lib.sclx_b200_last_error()
    rc = sclx.lib.sclx_b200_inclusive_scan(one.ctypes.data, one.ctypes.data, 1, sclx.I32, sclx.PLUS,
                                           None, 2, None, None)
    assert rc == sclx.E_INVALID_ARGUMENT
    rc = sclx.lib.sclx_b200_reduce_last_dim(None, 4, 4, sclx.F32, sclx.PLUS, one.ctypes.data, None, None)
    assert rc == sclx.E_INVALID_ARGUMENT
    rc = sclx.lib.sclx_b200_fill(one.ctypes.data, 1, 3, one.ctypes.data, None)
    assert rc in (sclx.E_UNSUPPORTED,)
    rc = sclx.lib.sclx_b200_reduce_last_dim_host(one.ctypes.data, 1, 1, 42, sclx.PLUS, one.ctypes.data,
                                                 one.ctypes.data)
    assert rc == sclx.E_UNSUPPORTED


def test_no_cpu_fallback(sclx):
    """without a device a compute call must fail loudly, never compute on the host"""
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    a = np.arange(16, dtype=np.int32)
    out = np.full(16, -1, np.int32)
    rc = sclx.lib.sclx_b200_inclusive_scan(a.ctypes.data, out.ctypes.data, 16, sclx.I32, sclx.PLUS,
                                           None, 0, None, None)
    assert rc > 0  # a cudaError_t
    assert np.all(out == -1)
    with pytest.raises(sclx.ScalixError):
        sclx.DeviceBuffer(1024)


def test_product_never_imports_the_oracle():
    for root in ("scalix_b200", "include"):
        for dirpath, _, files in os.walk(os.path.join(REPO, root)):
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                    text = open(os.path.join(dirpath, f), errors="replace").read()
                    assert "liboracle" not in text and "import oracle" not in text, os.path.join(dirpath, f)


def test_cmake_project_defines_the_reference_target(tmp_path):
    """the `scalix` CMake target of the reference (/root/reference CMakeLists.txt:13-44) exists
    here under the same name with the same switches, links the runtime library and is
    B200-only; configure only (no compilation)"""
    import shutil
    import subprocess

    text = open(os.path.join(REPO, "CMakeLists.txt")).read()
    assert "add_library(scalix INTERFACE)" in text
    for needle in ("--extended-lambda", "--expt-relaxed-constexpr", "SCALIX_DEBUG_KERNEL_LAUNCH",
                   "SCALIX_EMULATE_MULTIDEVICE", "set(CMAKE_CUDA_ARCHITECTURES 100a)",
                   "target_link_libraries(scalix INTERFACE scalix_b200)"):
        assert needle in text, needle
    cmake = shutil.which("cmake")
    if cmake is None or shutil.which("nvcc") is None:
        pytest.skip("cmake or nvcc not on PATH")
    gen = ["-G", "Ninja"] if shutil.which("ninja") else []
    p = subprocess.run([cmake, "-S", REPO, "-B", str(tmp_path / "b"), *gen, "-DSCALIX_EMULATE_MULTIDEVICE=4",
                        "-DSCALIX_DEBUG_KERNEL_LAUNCH=ON"], capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stdout[-2000:] + p.stderr[-2000:]
    assert "SCALIX_EMULATE_MULTIDEVICE is set to 4" in p.stdout
