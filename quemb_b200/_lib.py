"""ctypes binding of ``libqemb_b200.so`` (C ABI: ``include/qemb_b200.h``).

This is the thin host layer that replaces the pybind11 module
``quemb.molbe._cpp.eri_sparse_DF`` of the reference.  It never falls back to a CPU
implementation: if the shared library is missing, or no sm_100 device is visible, the
calls raise.
"""

from __future__ import annotations

import ctypes
import os
import re
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
ROOT = HERE.parent
LIB_PATH = HERE / "libqemb_b200.so"
HEADER = ROOT / "include" / "qemb_b200.h"

c_i64 = ctypes.c_int64
c_int = ctypes.c_int
c_dbl = ctypes.c_double
c_p = ctypes.c_void_p

QEMB_OK, QEMB_EINVAL, QEMB_ENOMEM, QEMB_ECUDA, QEMB_ENODEV, QEMB_EINDEX = 0, -1, -2, -3, -4, -5


class QembError(RuntimeError):
    """CUDA / argument errors (the reference raises ``std::runtime_error`` -> RuntimeError)."""


class QembNoDevice(QembError):
    """No usable B200: there is no CPU fallback behind this API."""


def build(verbose: bool = False) -> None:
    """Compile ``libqemb_b200.so`` for sm_100a (nvcc cross-compiles without a GPU)."""
    subprocess.run(
        ["make", "-C", str(HERE / "csrc"), "-j8"],
        check=True,
        stdout=None if verbose else subprocess.DEVNULL,
    )


def declared_symbols() -> list[str]:
    """Every function ``include/qemb_b200.h`` declares (used by the CPU export test)."""
    text = HEADER.read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(qemb_[A-Za-z0-9_]+)\s*\(", text)))


_LIB: ctypes.CDLL | None = None

_SIGS = {
    # name: (restype, argtypes)
    "qemb_last_error": (ctypes.c_char_p, []),
    "qemb_version": (ctypes.c_char_p, []),
    "qemb_get_log_level": (c_int, []),
    "qemb_set_log_level": (c_int, [c_int]),
    "qemb_tensor_count_nan": (c_int, [c_p, ctypes.POINTER(c_i64)]),
    "qemb_fragment_jk_enqueue": (c_int, [c_p, c_p, c_i64, c_p, c_p, c_i64]),
    "qemb_output_ring_depth": (c_int, []),
    "qemb_host_bandwidth": (c_int, [c_i64, ctypes.POINTER(c_dbl), ctypes.POINTER(c_dbl)]),
    "qemb_ctx_stream_handoff": (c_int, [c_p, c_p, c_int]),
    "qemb_ctx_autotune_output": (c_int, [c_p, ctypes.POINTER(c_dbl), ctypes.POINTER(c_dbl), ctypes.POINTER(c_dbl)]),
    "qemb_ctx_autotune_copy_alone": (c_int, [c_p, ctypes.POINTER(c_dbl)]),
    "qemb_ctx_autotune_trials": (c_int, [c_p, c_p]),
    "qemb_test_gram_tiles": (c_i64, [c_int, c_p]),
    "qemb_test_gram_splits": (c_int, [c_int, c_int, c_int, ctypes.POINTER(c_int), ctypes.POINTER(c_int)]),
    "qemb_test_balance_tiles": (c_int, [c_i64, c_p, c_int, c_p]),
    "qemb_autotune_pick": (c_int, [c_p, c_p, c_dbl, c_dbl, c_dbl, ctypes.POINTER(c_int)]),
    "qemb_device_count": (c_int, [ctypes.POINTER(c_int)]),
    "qemb_ctx_create": (c_int, [c_int, ctypes.POINTER(c_p)]),
    "qemb_ctx_destroy": (None, [c_p]),
    "qemb_ctx_synchronize": (c_int, [c_p]),
    "qemb_ctx_device": (c_int, [c_p, ctypes.POINTER(c_int)]),
    "qemb_ravel_symmetric": (c_i64, [c_i64, c_i64]),
    "qemb_unravel_symmetric": (None, [c_i64, ctypes.POINTER(c_i64), ctypes.POINTER(c_i64)]),
    "qemb_extract_unique": (c_i64, [c_i64, c_p, c_p, c_p, c_p]),
    "qemb_tensor_create": (c_int, [c_p, c_i64, c_i64, c_p, c_p, ctypes.POINTER(c_p)]),
    "qemb_tensor_destroy": (None, [c_p]),
    "qemb_tensor_n_unique": (c_int, [c_p, ctypes.POINTER(c_i64)]),
    "qemb_tensor_shape": (c_int, [c_p, ctypes.POINTER(c_i64), ctypes.POINTER(c_i64), ctypes.POINTER(c_i64)]),
    "qemb_tensor_column": (c_int, [c_p, c_i64, c_i64, ctypes.POINTER(c_i64)]),
    "qemb_tensor_reachable_columns": (c_int, [c_p, c_p]),
    "qemb_tensor_upload": (c_int, [c_p, c_p, c_i64, c_i64, c_i64]),
    "qemb_tensor_scatter_block": (c_int, [c_p, c_p, c_i64, c_i64, c_i64, c_i64]),
    "qemb_tensor_download": (c_int, [c_p, c_p, c_i64, c_i64, c_i64]),
    "qemb_metric_create": (c_int, [c_p, c_p, c_i64, c_i64, ctypes.POINTER(c_p)]),
    "qemb_metric_destroy": (None, [c_p]),
    "qemb_tensor_whiten": (c_int, [c_p, c_p]),
    "qemb_tensor_whiten_range": (c_int, [c_p, c_p, c_i64, c_i64]),
    "qemb_tensor_set_whitened": (c_int, [c_p, c_int]),
    "qemb_tensor_data_dev": (c_int, [c_p, ctypes.POINTER(c_p), ctypes.POINTER(c_i64)]),
    "qemb_tensor_is_whitened": (c_int, [c_p, ctypes.POINTER(c_int)]),
    "qemb_screen_create": (c_int, [c_p, c_p, c_i64, c_i64, ctypes.POINTER(c_p)]),
    "qemb_screen_destroy": (None, [c_p]),
    "qemb_workspace_create": (c_int, [c_p, ctypes.POINTER(c_p)]),
    "qemb_workspace_destroy": (None, [c_p]),
    "qemb_workspace_bytes": (c_int, [c_p, ctypes.POINTER(c_i64)]),
    "qemb_fragment_screen": (c_int, [c_p, c_p, c_p, c_i64, c_i64, c_i64, c_dbl, ctypes.POINTER(c_i64)]),
    "qemb_fragment_borderline": (c_int, [c_p, ctypes.POINTER(c_i64)]),
    "qemb_fragment_set_lists": (c_int, [c_p, c_p, c_i64, c_i64, c_i64, c_p, c_p]),
    "qemb_fragment_get_lists": (c_int, [c_p, c_p, c_p]),
    "qemb_fragment_contract_1st": (c_int, [c_p, c_p]),
    "qemb_fragment_contract_2nd": (c_int, [c_p]),
    "qemb_fragment_set_B": (c_int, [c_p, c_i64, c_p, c_i64]),
    "qemb_fragment_solve": (c_int, [c_p, c_p]),
    "qemb_fragment_gram": (c_int, [c_p]),
    "qemb_fragment_get_B": (c_int, [c_p, c_p, c_i64]),
    "qemb_fragment_get_sym": (c_int, [c_p, c_p, c_i64]),
    "qemb_fragment_get_G": (c_int, [c_p, c_p, c_i64]),
    "qemb_fragment_G_dev": (c_int, [c_p, ctypes.POINTER(c_p), ctypes.POINTER(c_i64)]),
    "qemb_fragment_transform": (
        c_int,
        [c_p, c_p, c_p, c_p, c_p, c_i64, c_i64, c_i64, c_dbl, c_i64, c_i64, c_p, c_i64, c_int, ctypes.POINTER(c_i64)],
    ),
    "qemb_fragment_wait_output": (c_int, [c_p, c_i64]),
    "qemb_fragment_jk": (c_int, [c_p, c_p, c_i64, c_p, c_p, c_i64]),
    "qemb_fragment_jk_time": (c_int, [c_p, ctypes.POINTER(ctypes.c_float), ctypes.POINTER(c_dbl)]),
    "qemb_tensor_copy_data": (c_int, [c_p, c_p]),
    "qemb_tensor_upload_whitened_rows": (c_int, [c_p, c_p, c_p, c_i64, c_i64, c_i64, c_i64]),
    "qemb_tensor_release_scratch": (c_int, [c_p]),
    "qemb_fragment_work": (c_int, [c_p, c_p, c_p]),
    "qemb_ctx_set_timing": (c_int, [c_p, c_int]),
    "qemb_ctx_set_output_mode": (c_int, [c_p, c_int]),
    "qemb_ctx_set_full_column_fraction": (c_int, [c_p, c_dbl]),
    "qemb_ctx_set_output_group_bytes": (c_int, [c_p, c_i64]),
    "qemb_ctx_set_k3_int8_slices": (c_int, [c_p, c_int]),
    "qemb_ctx_set_k3_int8_tolerance": (c_int, [c_p, c_dbl]),
    "qemb_k3_int8_bound": (c_int, [c_i64, c_int, ctypes.POINTER(c_dbl)]),
    "qemb_fragment_k3_choice": (c_int, [c_p, ctypes.POINTER(c_int), ctypes.POINTER(c_int)]),
    "qemb_k3_int8_slice_host": (c_int, [c_p, c_i64, c_i64, c_i64, c_int, c_p, c_p]),
    "qemb_host_threads": (c_int, [ctypes.POINTER(c_int)]),
    "qemb_host_mirror_lower": (c_int, [c_p, c_i64, c_i64]),
    "qemb_fragment_times": (c_int, [c_p, c_p]),
    "qemb_ctx_mark": (c_int, [c_p, c_int, c_int]),
    "qemb_ctx_mark_elapsed": (c_int, [c_p, c_int, c_int, ctypes.POINTER(ctypes.c_float)]),
    "qemb_ctx_launch_count": (c_int, [c_p, ctypes.POINTER(c_i64)]),
    "qemb_ctx_transfer_bytes": (c_int, [c_p, ctypes.POINTER(c_i64), ctypes.POINTER(c_i64)]),
    "qemb_ipc_alloc": (c_int, [c_p, c_i64, ctypes.POINTER(c_p), c_p]),
    "qemb_ipc_open": (c_int, [c_p, c_p, ctypes.POINTER(c_p)]),
    "qemb_ipc_close": (c_int, [c_p, c_p]),
    "qemb_ipc_free": (c_int, [c_p, c_p]),
    "qemb_device_zero": (c_int, [c_p, c_p, c_i64]),
    "qemb_device_mirror_lower": (c_int, [c_p, c_p, c_i64, c_i64]),
    "qemb_device_to_host": (c_int, [c_p, c_p, c_p, c_i64]),
    "qemb_host_alloc": (c_int, [ctypes.POINTER(c_p), c_i64]),
    "qemb_host_free": (c_int, [c_p]),
    "qemb_probe_dmma_tflops": (c_int, [c_p, c_i64, c_i64, c_int, ctypes.POINTER(c_dbl)]),
    "qemb_gram_noalloc": (c_int, [c_p, c_i64, c_i64, c_i64, c_p, c_i64, c_p]),
    "qemb_gram_noalloc_limits": (c_int, [ctypes.POINTER(c_i64), ctypes.POINTER(c_i64)]),
}


def lib() -> ctypes.CDLL:
    """Load the shared library (no compute happens here). Raises if it has not been built."""
    global _LIB
    if _LIB is None:
        path = Path(os.environ.get("QEMB_B200_LIB", LIB_PATH))
        if not path.exists():
            raise QembError(
                f"{path} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "or `make -C quemb_b200/csrc` (there is no CPU fallback)"
            )
        L = ctypes.CDLL(str(path))
        for name, (res, args) in _SIGS.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _LIB = L
    return _LIB


def last_error() -> str:
    return lib().qemb_last_error().decode()


def check(rc: int) -> None:
    if rc == QEMB_OK:
        return
    msg = last_error()
    if rc == QEMB_ENOMEM:
        raise MemoryError(msg)  # std::bad_alloc of the reference (cpp:56-62)
    if rc == QEMB_EINDEX:
        raise IndexError(msg)  # std::out_of_range (cpp:210)
    if rc == QEMB_ENODEV:
        raise QembNoDevice(msg)
    if rc == QEMB_EINVAL:
        raise QembError("invalid argument: " + msg)
    raise QembError(msg)


def ptr(a: np.ndarray | None) -> c_p:
    return c_p(None) if a is None else c_p(a.ctypes.data)


def device_count() -> int:
    n = c_int(0)
    rc = lib().qemb_device_count(ctypes.byref(n))
    if rc == QEMB_ENODEV:
        return 0
    check(rc)
    return n.value


def pinned_empty(shape, dtype=np.float64) -> np.ndarray:
    """A numpy array in page-locked host memory (asynchronous device -> host copies)."""
    dtype = np.dtype(dtype)
    n = int(np.prod(shape)) * dtype.itemsize
    p = c_p()
    check(lib().qemb_host_alloc(ctypes.byref(p), c_i64(max(n, 1))))
    buf = (ctypes.c_char * max(n, 1)).from_address(p.value)
    arr = np.frombuffer(buf, dtype=dtype, count=int(np.prod(shape))).reshape(shape)
    _PINNED[arr.ctypes.data] = p.value
    return arr


_PINNED: dict[int, int] = {}


def pinned_free(arr: np.ndarray) -> None:
    p = _PINNED.pop(arr.ctypes.data, None)
    if p is not None:
        check(lib().qemb_host_free(c_p(p)))
