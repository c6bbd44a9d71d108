"""Host-side mirror of the reference's native module ``quemb.molbe._cpp.eri_sparse_DF``
(API surface: ``src/quemb/molbe/_cpp/eri_sparse_DF.pyi``), backed by ``libqemb_b200.so``.

Same names, argument meaning and error behaviour as the reference, so that
``_run_sparse_df_driver(..., build_lowtri_PQ=GPU_MatrixHandle,
transform_integral_impl=transform_integral_cuda)`` (``src/quemb/molbe/eri_sparse_DF.py:535-656``)
works unchanged with this module in place of the Eigen extension.  All arithmetic runs on the
B200; nothing here computes on the CPU (index bookkeeping excepted).
"""

from __future__ import annotations

import ctypes
import logging
import weakref
from typing import Sequence

import numpy as np

from . import _lib
from ._lib import c_dbl, c_i64, c_int, c_p, check, lib, ptr
from .indexing import to_csr

__all__ = [
    "GPU_MatrixHandle",
    "SemiSparse3DTensor",
    "SemiSparseSym3DTensor",
    "contract_with_TA_1st",
    "contract_with_TA_2nd_to_sym_dense",
    "extract_unique",
    "get_AO_per_MO",
    "get_AO_reachable_by_MO_with_offset",
    "get_log_level",
    "set_log_level",
    "transform_integral",
    "transform_integral_cuda",
]

logger = logging.getLogger(__name__)

def set_log_level(log_level: int) -> None:
    """``indexers.hpp:43-46`` / ``pyi:139-157``: process-global level with Python's logging values (the reference
    driver forwards the root logger's level, ``eri_sparse_DF.py:605``).  At ``<= logging.INFO`` every
    ``transform_integral`` prints the reference's ``[TIMER] <step> finished in <t> s`` lines (``indexers.hpp:170-204``)
    from CUDA-event times of screening, K1, K2a, K2b and K3."""
    check(lib().qemb_set_log_level(c_int(int(log_level))))


def get_log_level() -> int:
    """``indexers.hpp:37-40``; default ``logging.ERROR`` like the reference's ``LOG_LEVEL``."""
    return int(lib().qemb_get_log_level())


def _f64(a, name: str) -> np.ndarray:
    a = np.asarray(a)
    if a.ndim != 2:
        raise _lib.QembError(f"{name} must be a matrix")
    return np.asfortranarray(a, dtype=np.float64)


# ---------------------------------------------------------------------------
# device context (one per device, created lazily)
# ---------------------------------------------------------------------------


class Context:
    """One CUDA device + its streams (``qemb_ctx``)."""

    def __init__(self, device: int = 0):
        h = c_p()
        check(lib().qemb_ctx_create(c_int(device), ctypes.byref(h)))
        self._h = h
        self.device = device
        self._finalizer = weakref.finalize(self, lib().qemb_ctx_destroy, h)

    def synchronize(self) -> None:
        check(lib().qemb_ctx_synchronize(self._h))

    def set_output_mode(self, triangle: bool) -> None:
        """True (default): host-bound G crosses PCIe as its lower triangle and is mirrored by host threads;
        False: the full square crosses PCIe."""
        check(lib().qemb_ctx_set_output_mode(self._h, c_int(1 if triangle else 0)))

    def set_full_column_fraction(self, frac: float) -> None:
        """Triangle output mode: the last ``frac`` of the columns travel whole over PCIe instead of being mirrored
        by the host threads."""
        check(lib().qemb_ctx_set_full_column_fraction(self._h, c_dbl(frac)))

    def set_output_group_bytes(self, nbytes: int) -> None:
        """Triangle output mode: bytes per copy group (the host mirrors a group once it has landed; default 256 MB)."""
        check(lib().qemb_ctx_set_output_group_bytes(self._h, c_i64(int(nbytes))))

    def autotune_output(self) -> dict[str, float]:
        """Measure this context's device->host copy rate alone and under the host mirror's contention plus the mirror
        rate, and set output mode / full-column fraction to the cheaper of full squares and a balanced split
        (``qemb_ctx_autotune_output``).
        With several ranks per node call it on all of them at the same moment (after a barrier)."""
        p, m, f, p0 = c_dbl(), c_dbl(), c_dbl(), c_dbl()
        check(lib().qemb_ctx_autotune_output(self._h, ctypes.byref(p), ctypes.byref(m), ctypes.byref(f)))
        check(lib().qemb_ctx_autotune_copy_alone(self._h, ctypes.byref(p0)))
        t = (c_dbl * 3)()
        check(lib().qemb_ctx_autotune_trials(self._h, t))
        trials = [float(x) for x in t]
        # which candidate the chosen mode is (0 full squares, 1 pure triangle, 2 the model's split) and whether the trials
        # alone decided: candidates within 5 % of the fastest trial are a tie that the rate model breaks
        chosen = 0 if f.value >= 1.0 else (1 if f.value <= 0.0 else 2)
        ran = [k for k, x in enumerate(trials) if x > 0]
        fastest = min(ran, key=lambda k: trials[k]) if ran else None
        return {"d2h_alone_gbs": p0.value, "d2h_gbs": p.value, "mirror_gbs": m.value, "upper_fraction_over_pcie": f.value,
                "trial_s_full_triangle_split": trials,
                "decided_by": "model" if fastest is None else ("trials" if fastest == chosen else "model (trials within 5 %)")}

    def set_k3_int8_slices(self, slices: int) -> None:
        """Opt-in K3 on the int8 tensor cores (0 = FP64 path, the default and the parity reference; 5 .. 8 integer
        slices of 8 bits: about 1e-11 / 2e-13 / FP64 rounding level (7, 8) of max|(ij|kl)| from the exact product, cost
        ~ slices (slices + 1) / 2; see ``profiles/k3_int8_parity_study_r1.txt``)."""
        check(lib().qemb_ctx_set_k3_int8_slices(self._h, c_int(slices)))

    def set_k3_int8_tolerance(self, abs_tol: float) -> None:
        """Guaranteed-tolerance mode of the opt-in int8 K3 (``qemb_ctx_set_k3_int8_tolerance``): per fragment the
        smallest slice count whose worst-case bound meets ``abs_tol`` on every (ij|kl), the FP64 kernel when none
        does; 0 switches it off."""
        check(lib().qemb_ctx_set_k3_int8_tolerance(self._h, c_dbl(abs_tol)))

    def set_timing(self, enabled: bool) -> None:
        check(lib().qemb_ctx_set_timing(self._h, c_int(1 if enabled else 0)))

    def mark(self, id: int, stream: str = "compute") -> None:
        check(lib().qemb_ctx_mark(self._h, c_int(id), c_int({"compute": 0, "prep": 1, "copy": 2}[stream])))

    def mark_elapsed_ms(self, id0: int, id1: int) -> float:
        ms = ctypes.c_float()
        check(lib().qemb_ctx_mark_elapsed(self._h, c_int(id0), c_int(id1), ctypes.byref(ms)))
        return float(ms.value)

    @property
    def launch_count(self) -> int:
        n = c_i64()
        check(lib().qemb_ctx_launch_count(self._h, ctypes.byref(n)))
        return n.value

    @property
    def transfer_bytes(self) -> tuple[int, int]:
        """(host->device, device->host) bytes this context has handed to the copy engines so far."""
        a, b = c_i64(), c_i64()
        check(lib().qemb_ctx_transfer_bytes(self._h, ctypes.byref(a), ctypes.byref(b)))
        return a.value, b.value

    def stream_handoff(self, external_stream: int, direction: int = 0) -> None:
        """Order an external CUDA stream (``torch.cuda.current_stream().cuda_stream``) against the compute stream
        without a host synchronisation: ``direction=0`` the external stream waits for the library's enqueued work,
        ``direction=1`` the library waits for the external stream (``qemb_ctx_stream_handoff``)."""
        check(lib().qemb_ctx_stream_handoff(self._h, c_p(int(external_stream)), c_int(direction)))

    def probe_dmma_tflops(self, n: int = 4096, k: int = 2048, iters: int = 5) -> float:
        v = c_dbl()
        check(lib().qemb_probe_dmma_tflops(self._h, c_i64(n), c_i64(k), c_int(iters), ctypes.byref(v)))
        return v.value


def host_bandwidth(nbytes: int = 1 << 30) -> tuple[float, float]:
    """(write GB/s, copy GB/s) of host DRAM as the library's host threads see it (``qemb_host_bandwidth``)."""
    w, c = c_dbl(), c_dbl()
    check(lib().qemb_host_bandwidth(c_i64(int(nbytes)), ctypes.byref(w), ctypes.byref(c)))
    return float(w.value), float(c.value)


def k3_int8_bound(naux: int, slices: int) -> float:
    """Worst case of ``|G_int8 - X^T X|`` per unit of ``2^(e_i + e_j)`` (``qemb_k3_int8_bound``)."""
    b = c_dbl()
    check(lib().qemb_k3_int8_bound(c_i64(naux), c_int(slices), ctypes.byref(b)))
    return float(b.value)


def output_ring_depth() -> int:
    """Host-output transforms that may be pending per workspace (``qemb_output_ring_depth``)."""
    return int(lib().qemb_output_ring_depth())


_CTX: dict[int, Context] = {}
_DEFAULT_DEVICE = 0


def set_default_device(device: int) -> None:
    global _DEFAULT_DEVICE
    _DEFAULT_DEVICE = int(device)


def default_context(device: int | None = None) -> Context:
    d = _DEFAULT_DEVICE if device is None else int(device)
    if d not in _CTX:
        _CTX[d] = Context(d)
    return _CTX[d]


class Workspace:
    """Per-caller scratch (``qemb_workspace``): plan, (P|mu i), (P|ij) and G buffers."""

    def __init__(self, ctx: Context | None = None):
        self.ctx = ctx or default_context()
        h = c_p()
        check(lib().qemb_workspace_create(self.ctx._h, ctypes.byref(h)))
        self._h = h
        self.last_ticket = -1
        # identity of the K1 result the workspace currently holds (SemiSparse3DTensor._resident); every call that plans
        # another fragment here drops it
        self._k1_token: object | None = None
        self._finalizer = weakref.finalize(self, lib().qemb_workspace_destroy, h)

    def wait_output(self, ticket: int) -> None:
        check(lib().qemb_fragment_wait_output(self._h, c_i64(ticket)))

    @property
    def device_bytes(self) -> int:
        n = c_i64()
        check(lib().qemb_workspace_bytes(self._h, ctypes.byref(n)))
        return n.value

    def times_ms(self) -> dict[str, float]:
        out = (ctypes.c_float * 5)()
        check(lib().qemb_fragment_times(self._h, out))
        return dict(zip(("screen", "K1", "K2a", "K2b", "K3"), [float(x) for x in out]))

    def work(self) -> dict[str, float]:
        out = np.zeros(8)
        check(lib().qemb_fragment_work(self._h, None, ptr(out)))
        keys = ["F_K1", "F_K2a", "F_K2b", "F_K3", "B_K1", "B_K2a", "B_K2b", "B_K3"]
        return dict(zip(keys, out.tolist()))

    def jk(self, dm: np.ndarray) -> tuple[np.ndarray, np.ndarray]:
        """``scf.hf.dot_eri_dm(eri, dm, hermi=1)`` on the (ij|kl) the last transform left on the device:
        ``vj[i,j] = sum_kl (ij|kl) dm[k,l]``, ``vk[i,k] = sum_jl (ij|kl) dm[j,l]`` (``qemb_fragment_jk``)."""
        D = _f64(dm, "dm")
        if D.shape[0] != D.shape[1]:
            raise _lib.QembError("dm must be square")
        n = D.shape[0]
        vj, vk = np.empty((n, n), order="F"), np.empty((n, n), order="F")
        check(lib().qemb_fragment_jk(self._h, ptr(D), c_i64(_ld(D)), ptr(vj), ptr(vk), c_i64(n)))
        return vj, vk

    def jk_enqueue(self, dm: np.ndarray, vj: np.ndarray, vk: np.ndarray) -> None:
        """:meth:`jk` enqueued on the compute stream (``qemb_fragment_jk_enqueue``): ``vj`` / ``vk`` are caller-owned
        ``(nmo, nmo)`` Fortran-ordered (ideally pinned) arrays that are complete after ``ctx.synchronize()``."""
        D = _f64(dm, "dm")
        n = D.shape[0]
        if D.shape[0] != D.shape[1]:
            raise _lib.QembError("dm must be square")
        for a in (vj, vk):
            if a.shape != (n, n) or a.dtype != np.float64 or not a.flags.f_contiguous:
                raise _lib.QembError("vj / vk must be (nmo, nmo) float64 Fortran-ordered arrays")
        check(lib().qemb_fragment_jk_enqueue(self._h, ptr(D), c_i64(_ld(D)), ptr(vj), ptr(vk), c_i64(n)))

    def jk_time(self) -> tuple[float, float]:
        """(device ms, algorithmic bytes) of the last :meth:`jk`."""
        ms, b = ctypes.c_float(), c_dbl()
        check(lib().qemb_fragment_jk_time(self._h, ctypes.byref(ms), ctypes.byref(b)))
        return float(ms.value), float(b.value)

    def G(self, npair: int) -> np.ndarray:
        """Download the resident (ij|kl) of the last transform."""
        out = np.empty((npair, npair), order="F")
        check(lib().qemb_fragment_get_G(self._h, ptr(out), c_i64(npair)))
        return out

    @property
    def k3_choice(self) -> tuple[int, int]:
        """(slices, max column exponent) the last K3 ran with; slices 0 = the FP64 kernel."""
        a, b = c_int(), c_int()
        check(lib().qemb_fragment_k3_choice(self._h, ctypes.byref(a), ctypes.byref(b)))
        return a.value, b.value

    @property
    def borderline(self) -> int:
        """Entries of the last screening within 1e-12 of ``MO_coeff_epsilon``."""
        n = c_i64()
        check(lib().qemb_fragment_borderline(self._h, ctypes.byref(n)))
        return n.value

    def lists(self, nmo: int) -> list[list[int]]:
        aptr = np.zeros(nmo + 1, dtype=np.int64)
        check(lib().qemb_fragment_get_lists(self._h, ptr(aptr), None))
        aidx = np.zeros(max(int(aptr[-1]), 1), dtype=np.int64)
        check(lib().qemb_fragment_get_lists(self._h, ptr(aptr), ptr(aidx)))
        return [aidx[aptr[i] : aptr[i + 1]].tolist() for i in range(nmo)]


_WS: dict[int, Workspace] = {}


def default_workspace(ctx: Context) -> Workspace:
    if ctx.device not in _WS:
        _WS[ctx.device] = Workspace(ctx)
    return _WS[ctx.device]


# ---------------------------------------------------------------------------
# GPU_MatrixHandle (cpp:64-107)
# ---------------------------------------------------------------------------


class GPU_MatrixHandle:
    """Lower Cholesky factor of (P|Q) resident on the device (``pyi:29-36``).

    Besides the copy of ``L`` the handle holds the inverted 128x128 diagonal blocks that the
    blocked tensor-core triangular solve needs.
    """

    def __init__(self, L_host: np.ndarray, ctx: Context | None = None) -> None:
        L = _f64(L_host, "L_host")
        if L.shape[0] != L.shape[1]:
            raise _lib.QembError("L_host must be square")
        self.ctx = ctx or default_context()
        self.shape = L.shape
        h = c_p()
        check(lib().qemb_metric_create(self.ctx._h, ptr(L), c_i64(L.strides[1] // 8), c_i64(L.shape[0]), ctypes.byref(h)))
        self._h = h
        self._finalizer = weakref.finalize(self, lib().qemb_metric_destroy, h)

    def __repr__(self) -> str:
        return f"GPU_MatrixHandle(rows={self.shape[0]}, cols={self.shape[1]}, device={self.ctx.device})"


class _ScreenHandle:
    def __init__(self, S_abs: np.ndarray, ctx: Context):
        S = _f64(S_abs, "S_abs")
        if S.shape[0] != S.shape[1]:
            raise _lib.QembError("S_abs must be square")
        self.ctx = ctx
        self.nao = S.shape[0]
        h = c_p()
        check(lib().qemb_screen_create(ctx._h, ptr(S), c_i64(S.strides[1] // 8), c_i64(S.shape[0]), ctypes.byref(h)))
        self._h = h
        self._finalizer = weakref.finalize(self, lib().qemb_screen_destroy, h)


# The reference passes the same S_abs / L_PQ ndarrays for every fragment (and pybind11 re-copies
# them each time, SURVEY 8b "Ownership").  Here they are uploaded once and found again by identity.
_SCREEN_CACHE: dict[tuple, tuple] = {}
_METRIC_CACHE: dict[tuple, tuple] = {}


def _cache_key(a: np.ndarray, ctx: Context) -> tuple:
    return (id(a), a.ctypes.data, a.shape, a.strides, ctx.device)


def _screen_for(S_abs, ctx: Context) -> _ScreenHandle:
    if isinstance(S_abs, _ScreenHandle):
        return S_abs
    a = np.asarray(S_abs)
    key = _cache_key(a, ctx)
    hit = _SCREEN_CACHE.get(key)
    if hit is not None and hit[0]() is a:
        return hit[1]
    h = _ScreenHandle(a, ctx)
    try:
        _SCREEN_CACHE.clear()  # one S_abs per molecule is live at a time
        _SCREEN_CACHE[key] = (weakref.ref(a), h)
    except TypeError:
        pass
    return h


def _metric_for(L_PQ, ctx: Context) -> GPU_MatrixHandle:
    if isinstance(L_PQ, GPU_MatrixHandle):
        return L_PQ
    a = np.asarray(L_PQ)
    key = _cache_key(a, ctx)
    hit = _METRIC_CACHE.get(key)
    if hit is not None and hit[0]() is a:
        return hit[1]
    h = GPU_MatrixHandle(a, ctx)
    try:
        _METRIC_CACHE.clear()
        _METRIC_CACHE[key] = (weakref.ref(a), h)
    except TypeError:
        pass
    return h


# ---------------------------------------------------------------------------
# SemiSparseSym3DTensor (cpp:110-298)
# ---------------------------------------------------------------------------


class SemiSparseSym3DTensor:
    """Screened, symmetric (P | mu nu) resident in HBM (``pyi:77-124``).

    ``SemiSparseSym3DTensor(shape, exch_reachable)`` allocates a NaN-filled tensor that is
    filled through ``mut_unique_dense_data`` (``eri_sparse_DF.py:477-488``);
    ``SemiSparseSym3DTensor(unique_dense_data, shape, exch_reachable)`` uploads the data.
    The 7-argument expert constructor of the reference is accepted; the redundant index
    structures it passes are rebuilt (bit-identically) on the host side of the library.
    """

    def __init__(self, *args, ctx: Context | None = None) -> None:
        data = None
        if len(args) == 2:
            shape, exch = args
        elif len(args) in (3, 7):
            data, shape, exch = args[:3]
        else:
            raise TypeError("SemiSparseSym3DTensor(shape, exch_reachable) or (unique_dense_data, shape, exch_reachable)")
        naux, nao, nao2 = (int(x) for x in shape)
        if nao != nao2 or len(exch) != nao:
            raise _lib.QembError("shape does not match exch_reachable")  # cpp:228-238
        self.ctx = ctx or default_context()
        self._shape = (naux, nao, nao)
        self._exch = [np.asarray(x, dtype=np.int64).reshape(-1) for x in exch]
        eptr, eidx = to_csr(self._exch)
        self._eptr, self._eidx = eptr, eidx
        h = c_p()
        check(lib().qemb_tensor_create(self.ctx._h, c_i64(naux), c_i64(nao), ptr(eptr), ptr(eidx), ctypes.byref(h)))
        self._h = h
        self._finalizer = weakref.finalize(self, lib().qemb_tensor_destroy, h)
        n = c_i64()
        check(lib().qemb_tensor_n_unique(h, ctypes.byref(n)))
        self._n_unique = n.value
        self._host_mirror: np.ndarray | None = None
        self._dirty = False
        self._has_device_data = False
        if data is not None:
            d = _f64(data, "unique_dense_data")
            if d.shape != (naux, self._n_unique):
                raise _lib.QembError(
                    f"unique_dense_data has shape {d.shape}, expected {(naux, self._n_unique)}"
                )
            self.upload(d)

    # -- data movement -------------------------------------------------------
    def upload(self, data: np.ndarray, col0: int = 0) -> None:
        d = _f64(data, "data")
        check(lib().qemb_tensor_upload(self._h, ptr(d), c_i64(d.strides[1] // 8 if d.shape[1] > 1 else d.shape[0]),
                                       c_i64(col0), c_i64(d.shape[1])))
        self._has_device_data = True
        if d is not self._host_mirror:
            self._device_changed()

    def set_whitened(self, flag: bool) -> None:
        """Declare the device data whitened / raw (``qemb_tensor_set_whitened``): before re-uploading raw columns into
        a tensor that was whitened, or after a distributed whitening."""
        check(lib().qemb_tensor_set_whitened(self._h, c_int(1 if flag else 0)))

    def scatter_block(self, integrals: np.ndarray, p0: int, q0: int) -> None:
        """Device-side replacement of the Python scatter loop of ``get_sparse_P_mu_nu``
        (``eri_sparse_DF.py:459-488``): ``integrals[i, j, :]`` (the C-ordered ``(n_i, n_j, naux)`` block
        ``_aux_e2`` returns for the AOs ``p0 + i`` / ``q0 + j``) goes to the column of every screened pair with
        ``q0 + j <= p0 + i``; one upload and one kernel per block instead of one numpy slice per pair."""
        self._flush()
        a = np.ascontiguousarray(integrals, dtype=np.float64)
        if a.ndim != 3 or a.shape[2] != self._shape[0]:
            raise _lib.QembError("integrals must have shape (n_i, n_j, naux)")
        check(lib().qemb_tensor_scatter_block(self._h, ptr(a), c_i64(a.shape[0]), c_i64(a.shape[1]), c_i64(p0), c_i64(q0)))
        self._has_device_data = True
        self._device_changed()

    def _flush(self) -> None:
        """Make the device copy current.  The host mirror handed out by :attr:`mut_unique_dense_data` stays alive
        (the reference returns a live view, ``cpp:810-812``): later accesses return the same array and later
        flushes upload it again."""
        if self._dirty and self._host_mirror is not None:
            self._dirty = False
            # a view handed out from a whitened tensor holds whitened values: writing them back must not turn the
            # flag off (a full upload clears it)
            was_whitened = self._has_device_data and self.whitened
            self.upload(self._host_mirror)
            if was_whitened:
                self.set_whitened(True)

    def _device_changed(self) -> None:
        """The device data were modified behind the host mirror (whiten, block staging, upload): the next access to
        :attr:`mut_unique_dense_data` downloads a fresh copy instead of resurrecting stale (or NaN) values."""
        self._host_mirror = None
        self._dirty = False

    def release_host_mirror(self) -> None:
        """Drop the host staging copy (after a final flush); the device copy is the tensor from here on."""
        self._flush()
        self._device_changed()

    @property
    def mut_unique_dense_data(self) -> np.ndarray:
        """Writable host view (``pyi:104-110``): NaN-filled like ``cpp:160`` for a fresh tensor, otherwise the
        current device content; writes reach the device at the next device use.  Re-read the property for every
        batch of writes (the reference driver does, ``eri_sparse_DF.py:486-488``): a view kept across a device use
        and written afterwards is only picked up at the next property access."""
        if self._host_mirror is None:
            if self._has_device_data:
                self._host_mirror = self._download()
            else:
                self._host_mirror = np.full((self._shape[0], self._n_unique), np.nan, order="F")
        self._dirty = True
        return self._host_mirror

    def _download(self) -> np.ndarray:
        out = np.empty((self._shape[0], self._n_unique), order="F")
        check(lib().qemb_tensor_download(self._h, ptr(out), c_i64(self._shape[0]), c_i64(0), c_i64(self._n_unique)))
        return out

    def count_nan(self) -> int:
        """NaN entries left on the device (``qemb_tensor_count_nan``; the reference's post-staging assert, ``py:492``)."""
        self._flush()
        n = c_i64()
        check(lib().qemb_tensor_count_nan(self._h, ctypes.byref(n)))
        return n.value

    @property
    def unique_dense_data(self) -> np.ndarray:
        self._flush()
        return self._download()

    def __getitem__(self, key: tuple[int, int]) -> np.ndarray:
        mu, nu = key
        col = c_i64()
        check(lib().qemb_tensor_column(self._h, c_i64(mu), c_i64(nu), ctypes.byref(col)))
        self._flush()
        out = np.empty((self._shape[0], 1), order="F")
        check(lib().qemb_tensor_download(self._h, ptr(out), c_i64(self._shape[0]), col, c_i64(1)))
        return out[:, 0]

    # -- metric pre-application (design option O1) ----------------------------
    def whiten(self, L_PQ) -> None:
        """Replace (P|mu nu) by L^-1 (P|mu nu) once; later transforms skip the per-fragment solve."""
        self._flush()
        m = _metric_for(L_PQ, self.ctx)
        check(lib().qemb_tensor_whiten(self._h, m._h))
        self._device_changed()

    @classmethod
    def whitened_slab(cls, unique_dense_data: np.ndarray, shape, exch_reachable, L_PQ, p_range: tuple[int, int],
                      *, ctx: Context | None = None, chunk_cols: int = 8192) -> "SemiSparseSym3DTensor":
        """Rows ``[p_begin, p_end)`` of ``L^-1 (P|mu nu)`` as a tensor of its own (aux-sharded fragments,
        SURVEY 8e): the host matrix is streamed through the device in column chunks, solved against the
        metric there, and only the slab's rows stay resident, so a rank never holds all of (P|mu nu).
        Transforms on the result (metric ``None``) give the slab's partial (ij|kl)."""
        naux, nao, _ = (int(x) for x in shape)
        p0, p1 = int(p_range[0]), int(p_range[1])
        if not (0 <= p0 < p1 <= naux) or p0 % 16:
            raise _lib.QembError("auxiliary slab must satisfy 0 <= p_begin < p_end <= naux, p_begin % 16 == 0")
        t = cls((p1 - p0, nao, nao), exch_reachable, ctx=ctx)
        d = np.asarray(unique_dense_data)
        if d.shape != (naux, t._n_unique) or d.dtype != np.float64 or d.strides[0] != 8:
            d = _f64(d, "unique_dense_data")
            if d.shape != (naux, t._n_unique):
                raise _lib.QembError(f"unique_dense_data has shape {d.shape}, expected {(naux, t._n_unique)}")
        m = _metric_for(L_PQ, t.ctx)
        ld = d.strides[1] // 8 if d.shape[1] > 1 else naux
        for c0 in range(0, t._n_unique, chunk_cols):
            nc = min(chunk_cols, t._n_unique - c0)
            check(lib().qemb_tensor_upload_whitened_rows(t._h, m._h, c_p(d.ctypes.data + 8 * ld * c0), c_i64(ld),
                                                         c_i64(c0), c_i64(nc), c_i64(p0)))
        check(lib().qemb_tensor_release_scratch(t._h))
        t._has_device_data = True
        t.p_range = (p0, p1)
        return t

    def whiten_distributed(self, L_PQ, dist, group=None) -> None:
        """``whiten`` for ranks that each hold a replica of (P|mu nu) (fragment-parallel builds): rank r solves only
        its 1/world share of the columns and the shares are exchanged in place with one NCCL broadcast per rank
        (NVLink), instead of every rank repeating all ``naux^2 n_unique`` flops."""
        import torch  # noqa: PLC0415

        self._flush()
        m = _metric_for(L_PQ, self.ctx)
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        bounds = [(self._n_unique * r) // world for r in range(world + 1)]
        check(lib().qemb_tensor_whiten_range(self._h, m._h, c_i64(bounds[rank]), c_i64(bounds[rank + 1] - bounds[rank])))
        dev, ld = c_p(), c_i64()
        check(lib().qemb_tensor_data_dev(self._h, ctypes.byref(dev), ctypes.byref(ld)))
        self.ctx.synchronize()  # the collective runs on torch's stream

        class _View:  # CUDA array interface over the library's device memory
            def __init__(self, ptr_, n):
                self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f8", "data": (ptr_, False), "version": 2}

        works = []
        for r in range(world):
            n = (bounds[r + 1] - bounds[r]) * ld.value
            if n == 0:
                continue
            view = torch.as_tensor(_View(dev.value + 8 * bounds[r] * ld.value, n), device=torch.device("cuda", self.ctx.device))
            src = r if group is None else dist.get_global_rank(group, r)
            works.append(dist.broadcast(view, src=src, group=group, async_op=True))
        for w in works:
            w.wait()
        torch.cuda.current_stream(self.ctx.device).synchronize()
        check(lib().qemb_tensor_set_whitened(self._h, c_int(1)))
        self._has_device_data = True
        self._device_changed()

    def copy_data_from(self, other: "SemiSparseSym3DTensor") -> None:
        """Device-to-device copy of the data of a tensor built from the same lists."""
        other._flush()
        check(lib().qemb_tensor_copy_data(self._h, other._h))
        self._has_device_data = other._has_device_data
        self._device_changed()

    @property
    def whitened(self) -> bool:
        f = c_int()
        check(lib().qemb_tensor_is_whitened(self._h, ctypes.byref(f)))
        return bool(f.value)

    def __repr__(self) -> str:  # cpp:214-225
        naux, nao1, nao2 = self._shape
        nz, size = self.nonzero_size, self.size
        return (f"Semi-sparse symmetric 3D tensor, shape=({naux}, {nao1}, {nao2}), n non-zero, unique elements = {nz}, "
                f"n dense elements = {size},\ncompression (compared to non-symmetric, dense storage) = "
                f"{100.0 * nz / max(size, 1):g}% \n")

    # -- index structures -------------------------------------------------------
    @property
    def shape(self) -> tuple[int, int, int]:
        return self._shape

    @property
    def size(self) -> int:
        return self._shape[0] * self._shape[1] * self._shape[2]

    @property
    def nonzero_size(self) -> int:
        return self._shape[0] * self._n_unique

    @property
    def exch_reachable(self) -> list[list[int]]:
        return [x.tolist() for x in self._exch]

    @property
    def exch_reachable_unique(self) -> list[list[int]]:
        return extract_unique(self._exch)

    @property
    def offsets(self) -> dict[int, int]:
        out: dict[int, int] = {}
        counter = 0
        for mu, lst in enumerate(self.exch_reachable_unique):
            for nu in lst:
                out[int(lib().qemb_ravel_symmetric(c_i64(mu), c_i64(nu)))] = counter
                counter += 1
        return out

    @property
    def exch_reachable_with_offsets(self) -> list[list[tuple[int, int]]]:
        col = np.zeros(max(self._eidx.size, 1), dtype=np.int64)
        check(lib().qemb_tensor_reachable_columns(self._h, ptr(col)))
        return [
            [(int(col[k]), int(self._eidx[k])) for k in range(self._eptr[mu], self._eptr[mu + 1])]
            for mu in range(self._shape[1])
        ]


# ---------------------------------------------------------------------------
# SemiSparse3DTensor (cpp:300-437): result of K1
# ---------------------------------------------------------------------------


class SemiSparse3DTensor:
    """(P | mu i), the result of the first contraction (``cpp:300-437``, ``pyi:38-75``).

    Two kinds of instances behave alike:

    * what :func:`contract_with_TA_1st` returns: the data live in the workspace that produced them (device
      resident) and :attr:`dense_data` downloads them;
    * what the reference's three public constructors build from host data (``cpp:303-347``, ``pyi:41-58``):
      ``SemiSparse3DTensor(shape, AO_reachable_by_MO)`` (NaN-filled like ``cpp:346``),
      ``SemiSparse3DTensor(dense_data, shape, AO_reachable_by_MO)`` and the 5-argument expert form; the data stay on
      the host until :func:`contract_with_TA_2nd_to_sym_dense` uploads them (``qemb_fragment_set_B``).

    ``mut_dense_data`` is a writable host view in both cases; a K1 result that was written through it is treated like
    host data from then on.
    """

    def __init__(self, *args, ctx: Context | None = None) -> None:
        data = with_off = offsets = None
        if len(args) == 2:
            shape, lists = args
        elif len(args) == 3:
            data, shape, lists = args
        elif len(args) == 5:
            data, shape, lists, with_off, offsets = args
        else:
            raise TypeError("SemiSparse3DTensor(shape, AO_reachable_by_MO), (dense_data, shape, AO_reachable_by_MO) or "
                            "(dense_data, shape, AO_reachable_by_MO, AO_reachable_by_MO_with_offsets, offsets)")
        self._shape = tuple(int(x) for x in shape)
        if len(self._shape) != 3:
            raise TypeError("shape must be (naux, nao, nmo)")
        self._lists = [list(map(int, x)) for x in lists]
        self._n = sum(len(x) for x in self._lists)
        self._ctx = ctx
        self._ws: Workspace | None = None
        self._token = None
        self._TA = None
        if with_off is None:
            # build_offsets_and_pairs, cpp:405-423: running counter over i then mu; NOTE the key of the offsets map
            # is ravel_Fortran(i_MO, mu, nao) = i + mu * nao here, while K1 files mu + i * nao (cpp:515)
            self._with_off = get_AO_reachable_by_MO_with_offset(self._lists)
            nao = self._shape[1]
            self._offsets = {i + mu * nao: off for i, row in enumerate(self._with_off) for off, mu in row}
        else:
            self._with_off = [[(int(o), int(mu)) for o, mu in row] for row in with_off]
            self._offsets = {int(k): int(v) for k, v in dict(offsets).items()}
        if data is None:
            self._host = np.full((self._shape[0], self._n), np.nan, order="F")  # cpp:346
        else:
            self._host = np.array(data, dtype=np.float64, order="F")  # pybind11 copies numpy -> Eigen
            if self._host.ndim != 2:
                raise TypeError("dense_data must be a matrix")

    @classmethod
    def _from_workspace(cls, ws: Workspace, shape, AO_reachable_by_MO, TA: np.ndarray) -> "SemiSparse3DTensor":
        """The K1 result left in ``ws`` (:func:`contract_with_TA_1st`)."""
        self = cls.__new__(cls)
        self._shape = tuple(int(x) for x in shape)
        self._lists = [list(map(int, x)) for x in AO_reachable_by_MO]
        self._n = sum(len(x) for x in self._lists)
        self._ctx = ws.ctx
        self._ws = ws
        self._token = ws._k1_token = object()
        self._TA = TA
        self._with_off = get_AO_reachable_by_MO_with_offset(self._lists)
        nao = self._shape[1]
        self._offsets = {mu + i * nao: off for i, row in enumerate(self._with_off) for off, mu in row}  # cpp:513-518
        self._host = None
        return self

    def _resident(self) -> bool:
        """True while the workspace still holds this tensor's columns."""
        return self._ws is not None and self._ws._k1_token is self._token

    def __repr__(self) -> str:  # cpp:386-398
        naux, nao, nmo = self._shape
        nz, size = self.nonzero_size, self.size
        return (f"Semi-sparse 3D tensor, shape=({naux}, {nao}, {nmo}), n non-zero, unique elements = {nz}, "
                f"n dense elements = {size},\ncompression (compared to dense storage) = {100.0 * nz / max(size, 1):g}% \n")

    @property
    def shape(self):
        return self._shape

    @property
    def size(self) -> int:
        return self._shape[0] * self._shape[1] * self._shape[2]

    @property
    def nonzero_size(self) -> int:
        if self._host is not None:
            return int(self._host.size)  # _dense_data.size(), cpp:381
        return self._shape[0] * self._n

    @property
    def AO_reachable_by_MO(self) -> list[list[int]]:
        return self._lists

    @property
    def AO_reachable_by_MO_with_offsets(self):
        return self._with_off

    @property
    def offsets(self) -> dict[int, int]:
        return self._offsets

    def _download(self) -> np.ndarray:
        if not self._resident():
            raise _lib.QembError("the workspace that held this (P | mu i) has been reused for another fragment; "
                                 "read dense_data before the next fragment, or give each tensor its own Workspace")
        out = np.zeros((self._shape[0], max(self._n, 1)), order="F")
        if self._n:
            check(lib().qemb_fragment_get_B(self._ws._h, ptr(out), c_i64(self._shape[0])))
        return out[:, : self._n]

    @property
    def dense_data(self) -> np.ndarray:
        if self._host is not None:
            v = self._host.view()
            v.flags.writeable = False
            return v
        return self._download()

    @property
    def mut_dense_data(self) -> np.ndarray:
        """Writable view (``pyi:66-67``, ``cpp:363-366``).  For a K1 result the columns are brought to the host first
        and the tensor is host data from then on (the next K2a uploads whatever was written)."""
        if self._host is None:
            self._host = np.asfortranarray(self._download())
            self._ws = self._token = None
        return self._host

    def __getitem__(self, key):
        a, b = key
        k = int(a) + int(b) * self._shape[1]  # get_aux_vector: _offsets.at(ravel_Fortran(mu, i, nao)), cpp:383-387
        if k not in self._offsets:
            raise IndexError(f"({a}, {b}) is not stored in this tensor")  # std::out_of_range
        return self.dense_data[:, self._offsets[k]]


# ---------------------------------------------------------------------------
# free functions
# ---------------------------------------------------------------------------


def extract_unique(exch_reachable: Sequence[Sequence[int]]) -> list[list[int]]:
    """``indexers.hpp:149-162``."""
    eptr, eidx = to_csr(exch_reachable)
    nao = len(eptr) - 1
    uptr = np.zeros(nao + 1, dtype=np.int64)
    n = lib().qemb_extract_unique(c_i64(nao), ptr(eptr), ptr(eidx), ptr(uptr), None)
    if n < 0:
        check(int(n))
    uidx = np.zeros(max(int(n), 1), dtype=np.int64)
    lib().qemb_extract_unique(c_i64(nao), ptr(eptr), ptr(eidx), ptr(uptr), ptr(uidx))
    return [uidx[uptr[i] : uptr[i + 1]].tolist() for i in range(nao)]


def get_AO_reachable_by_MO_with_offset(AO_reachable_by_MO) -> list[list[tuple[int, int]]]:
    """``cpp:467-482``."""
    out, counter = [], 0
    for aos in AO_reachable_by_MO:
        row = []
        for ao in aos:
            row.append((counter, int(ao)))
            counter += 1
        out.append(row)
    return out


def get_AO_per_MO(TA, S_abs, epsilon: float, *, ctx: Context | None = None, ws: Workspace | None = None) -> list[list[int]]:
    """``cpp:443-465``: A(i) = {mu : |S_abs TA|(mu, i) >= epsilon}; evaluated on the device."""
    ctx = ctx or default_context()
    ws = ws or default_workspace(ctx)
    TA = _f64(TA, "TA")
    sc = _screen_for(S_abs, ctx)
    n = c_i64()
    ws._k1_token = None
    check(lib().qemb_fragment_screen(ws._h, sc._h, ptr(TA), c_i64(TA.strides[1] // 8 if TA.shape[1] > 1 else TA.shape[0]),
                                     c_i64(TA.shape[0]), c_i64(TA.shape[1]), c_dbl(epsilon), ctypes.byref(n)))
    return ws.lists(TA.shape[1])


def _ld(a: np.ndarray) -> int:
    return a.strides[1] // 8 if a.shape[1] > 1 else a.shape[0]


def contract_with_TA_1st(TA, int_P_mu_nu: SemiSparseSym3DTensor, AO_by_MO, *, ws: Workspace | None = None) -> SemiSparse3DTensor:
    """K1, ``cpp:484-532``."""
    ctx = int_P_mu_nu.ctx
    ws = ws or default_workspace(ctx)
    TA = _f64(TA, "TA")
    int_P_mu_nu._flush()
    aptr, aidx = to_csr(AO_by_MO)
    check(lib().qemb_fragment_set_lists(ws._h, ptr(TA), c_i64(_ld(TA)), c_i64(TA.shape[0]), c_i64(TA.shape[1]),
                                        ptr(aptr), ptr(aidx)))
    check(lib().qemb_fragment_contract_1st(ws._h, int_P_mu_nu._h))
    naux = int_P_mu_nu.shape[0]
    return SemiSparse3DTensor._from_workspace(ws, (naux, TA.shape[0], TA.shape[1]), AO_by_MO, TA)


def contract_with_TA_2nd_to_sym_dense(int_mu_i_P: SemiSparse3DTensor, TA, *, ws: Workspace | None = None) -> np.ndarray:
    """K2a, ``cpp:560-605``: (P | ij), pair column T(max)+min, lists of min / TA column of max.

    A tensor that :func:`contract_with_TA_1st` left on the device is contracted where it lives when ``TA`` is the
    matrix K1 ran with; a tensor built from host data (the reference's public constructors, ``mut_dense_data``), or a
    different ``TA``, goes through ``qemb_fragment_set_lists`` + ``qemb_fragment_set_B`` first."""
    naux, nao, nmo = int_mu_i_P.shape
    TA = _f64(TA, "TA")
    if TA.shape != (nao, nmo):
        raise _lib.QembError(f"TA has shape {TA.shape}, the tensor wants {(nao, nmo)}")  # assert cpp:576-577
    same_TA = int_mu_i_P._TA is not None and (TA is int_mu_i_P._TA or np.array_equal(TA, int_mu_i_P._TA))
    if int_mu_i_P._resident() and same_TA:
        ws = int_mu_i_P._ws
    else:
        host = int_mu_i_P.dense_data  # downloads a K1 result whose TA differs (raises if the workspace moved on)
        ws = ws or default_workspace(int_mu_i_P._ctx or default_context())
        with_off = int_mu_i_P.AO_reachable_by_MO_with_offsets
        lists = [[mu for _, mu in row] for row in with_off]
        cols = np.fromiter((off for row in with_off for off, _ in row), dtype=np.int64)
        if cols.size and (cols.min() < 0 or cols.max() >= host.shape[1]):
            raise IndexError("an offset of AO_reachable_by_MO_with_offsets lies outside dense_data")
        if host.shape[0] != naux:
            raise _lib.QembError("dense_data has the wrong number of rows")
        # the library's column order is the running counter over (i, mu) (cpp:467-482); the expert constructor may
        # carry any offsets
        B = host if np.array_equal(cols, np.arange(cols.size)) and host.shape[1] == cols.size else host[:, cols]
        B = np.asfortranarray(B)
        aptr, aidx = to_csr(lists)
        ws._k1_token = None
        check(lib().qemb_fragment_set_lists(ws._h, ptr(TA), c_i64(_ld(TA)), c_i64(nao), c_i64(nmo), ptr(aptr), ptr(aidx)))
        check(lib().qemb_fragment_set_B(ws._h, c_i64(naux), ptr(B) if B.size else None, c_i64(max(naux, 1))))
    check(lib().qemb_fragment_contract_2nd(ws._h))
    out = np.empty((naux, nmo * (nmo + 1) // 2), order="F")
    check(lib().qemb_fragment_get_sym(ws._h, ptr(out), c_i64(naux)))
    return out


def transform_integral(
    int_P_mu_nu: SemiSparseSym3DTensor,
    TA,
    S_abs,
    L_PQ,
    MO_coeff_epsilon: float,
    *,
    ws: Workspace | None = None,
    out: np.ndarray | None = None,
    synchronize: bool = True,
    p_range: tuple[int, int] = (0, -1),
) -> np.ndarray:
    """``transform_integral`` / ``transform_integral_cuda`` (``cpp:724-751``, ``pyi:158-181``).

    ``L_PQ`` may be the host matrix (as for the reference's CPU entry point) or a
    :class:`GPU_MatrixHandle`.  Returns the ``(npair, npair)`` matrix of (ij|kl) in s4 pair order.
    """
    ctx = int_P_mu_nu.ctx
    ws = ws or default_workspace(ctx)
    TA = _f64(TA, "TA")
    int_P_mu_nu._flush()
    sc = _screen_for(S_abs, ctx)
    m = None if (L_PQ is None) else _metric_for(L_PQ, ctx)
    nmo = TA.shape[1]
    npair = nmo * (nmo + 1) // 2
    if out is None:
        out = np.empty((npair, npair), order="F")
    if out.shape != (npair, npair) or out.dtype != np.float64:
        raise _lib.QembError("out has the wrong shape or dtype")
    ld_out = npair
    if not (out.flags.f_contiguous or out.flags.c_contiguous):
        raise _lib.QembError("out must be contiguous")  # G is symmetric: C and F order coincide
    ticket = c_i64()
    ws._k1_token = None
    check(lib().qemb_fragment_transform(
        ws._h, int_P_mu_nu._h, None if m is None else m._h, sc._h, ptr(TA), c_i64(_ld(TA)), c_i64(TA.shape[0]),
        c_i64(nmo), c_dbl(MO_coeff_epsilon), c_i64(p_range[0]), c_i64(p_range[1]), ptr(out), c_i64(ld_out), c_int(0),
        ctypes.byref(ticket)))
    ws.last_ticket = ticket.value
    if synchronize:
        ws.wait_output(ticket.value)  # copies AND the host-side mirror of the upper triangle
    return out


def transform_integral_device(
    int_P_mu_nu: SemiSparseSym3DTensor,
    TA,
    S_abs,
    L_PQ,
    MO_coeff_epsilon: float,
    out_dev_ptr: int,
    ld_out: int,
    *,
    ws: Workspace | None = None,
    p_range: tuple[int, int] = (0, -1),
) -> None:
    """Same path, but (ij|kl) is written straight into caller-owned device memory (e.g. a torch
    tensor's ``data_ptr()``) and nothing travels to the host.  Asynchronous."""
    ctx = int_P_mu_nu.ctx
    ws = ws or default_workspace(ctx)
    TA = _f64(TA, "TA")
    int_P_mu_nu._flush()
    sc = _screen_for(S_abs, ctx)
    m = None if (L_PQ is None) else _metric_for(L_PQ, ctx)
    ws._k1_token = None
    check(lib().qemb_fragment_transform(
        ws._h, int_P_mu_nu._h, None if m is None else m._h, sc._h, ptr(TA), c_i64(_ld(TA)), c_i64(TA.shape[0]),
        c_i64(TA.shape[1]), c_dbl(MO_coeff_epsilon), c_i64(p_range[0]), c_i64(p_range[1]), c_p(out_dev_ptr),
        c_i64(ld_out), c_int(1), None))


def transform_integral_resident(int_P_mu_nu: SemiSparseSym3DTensor, TA, S_abs, L_PQ, MO_coeff_epsilon: float, *,
                                ws: Workspace | None = None) -> Workspace:
    """Same path, (ij|kl) stays in the workspace (nothing travels to the host): the entry point of the
    fused consumers (``Workspace.jk``; ``quemb_b200.consumer``).  Returns the workspace that holds G."""
    ctx = int_P_mu_nu.ctx
    ws = ws or default_workspace(ctx)
    TA = _f64(TA, "TA")
    int_P_mu_nu._flush()
    sc = _screen_for(S_abs, ctx)
    m = None if (L_PQ is None) else _metric_for(L_PQ, ctx)
    ws._k1_token = None
    check(lib().qemb_fragment_transform(
        ws._h, int_P_mu_nu._h, None if m is None else m._h, sc._h, ptr(TA), c_i64(_ld(TA)), c_i64(TA.shape[0]),
        c_i64(TA.shape[1]), c_dbl(MO_coeff_epsilon), c_i64(0), c_i64(-1), None, c_i64(0), c_int(0), None))
    return ws


def transform_integral_cuda(int_P_mu_nu, TA, S_abs, L_PQ: GPU_MatrixHandle, MO_coeff_epsilon: float) -> np.ndarray:
    """Name of the reference's GPU entry point (``cpp:724-736``); same implementation."""
    return transform_integral(int_P_mu_nu, TA, S_abs, L_PQ, MO_coeff_epsilon)
