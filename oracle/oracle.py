"""oracle/oracle.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

ctypes front-end of the plain-C restatement ``oracle/eri_oracle.c`` plus a
loader for the reference's own compiled extension (``oracle/_ref``).  Only
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline /
``--impl reference`` leg may import this module; nothing under ``quemb_b200/``
does.

Parity status: PINNED against the compiled reference (see eri_oracle.c header,
``oracle/make_golden.py`` and ``tests/golden/``).
"""

from __future__ import annotations

import ctypes
import importlib.util
import os
import subprocess
import sys
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
_LIB = None

c_i64 = ctypes.c_int64
c_dbl = ctypes.c_double
_P = ctypes.c_void_p


def build() -> None:
    """Compile the C restatement and, if the reference sources are present, oracle/_ref."""
    subprocess.run(["make", "-C", str(HERE), "-j2", "all"], check=True, stdout=subprocess.DEVNULL)


def lib() -> ctypes.CDLL:
    global _LIB
    if _LIB is None:
        so = HERE / "liberi_oracle.so"
        if not so.exists():
            subprocess.run(["make", "-C", str(HERE), "liberi_oracle.so"], check=True,
                           stdout=subprocess.DEVNULL)
        _LIB = ctypes.CDLL(str(so))
        _LIB.orc_ravel_symmetric.restype = c_i64
        _LIB.orc_ravel_symmetric.argtypes = [c_i64, c_i64]
        _LIB.orc_extract_unique.restype = c_i64
    return _LIB


def _ptr(a: np.ndarray):
    return ctypes.c_void_p(a.ctypes.data)


def _f(a) -> np.ndarray:
    return np.asfortranarray(a, dtype=np.float64)


def _i(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.int64)


def to_csr(lists):
    n = len(lists)
    ptr = np.zeros(n + 1, dtype=np.int64)
    arrs = [np.asarray(x, dtype=np.int64).reshape(-1) for x in lists]
    if n:
        ptr[1:] = np.cumsum([a.size for a in arrs])
    idx = np.concatenate(arrs) if n and ptr[-1] > 0 else np.zeros(0, dtype=np.int64)
    return ptr, _i(idx)


def from_csr(ptr, idx):
    return [idx[ptr[i]: ptr[i + 1]].tolist() for i in range(len(ptr) - 1)]


# --- index maps -----------------------------------------------------------

def ravel_symmetric(a: int, b: int) -> int:
    return int(lib().orc_ravel_symmetric(a, b))


def unravel_symmetric(i: int) -> tuple[int, int]:
    out = np.zeros(2, dtype=np.int64)
    lib().orc_unravel_symmetric(c_i64(i), _ptr(out))
    return int(out[0]), int(out[1])


def extract_unique(exch_reachable) -> tuple[np.ndarray, np.ndarray]:
    ptr, idx = to_csr(exch_reachable)
    nao = len(ptr) - 1
    uptr = np.zeros(nao + 1, dtype=np.int64)
    n = lib().orc_extract_unique(c_i64(nao), _ptr(ptr), _ptr(idx), _ptr(uptr), None)
    uidx = np.zeros(max(int(n), 1), dtype=np.int64)
    lib().orc_extract_unique(c_i64(nao), _ptr(ptr), _ptr(idx), _ptr(uptr), _ptr(uidx))
    return uptr, uidx[: int(n)]


def reachable_columns(exch_reachable):
    """-> (rptr, ridx, rcol, uptr, uidx)"""
    rptr, ridx = to_csr(exch_reachable)
    uptr, uidx = extract_unique(exch_reachable)
    rcol = np.zeros(max(ridx.size, 1), dtype=np.int64)
    uidx_ = uidx if uidx.size else np.zeros(1, dtype=np.int64)
    lib().orc_reachable_columns(c_i64(len(rptr) - 1), _ptr(rptr), _ptr(ridx), _ptr(uptr), _ptr(uidx_), _ptr(rcol))
    return rptr, ridx, rcol[: ridx.size], uptr, uidx


# --- numerical steps --------------------------------------------------------

def get_AO_per_MO(TA, S_abs, eps: float, return_X: bool = False):
    TA, S_abs = _f(TA), _f(S_abs)
    nao, nmo = TA.shape
    mask = np.zeros((nao, nmo), dtype=np.uint8, order="F")
    X = np.zeros((nao, nmo), dtype=np.float64, order="F")
    lib().orc_screen_mask(c_i64(nao), c_i64(nmo), _ptr(S_abs), _ptr(TA), c_dbl(eps), _ptr(mask), _ptr(X))
    lists = [np.nonzero(mask[:, i])[0].astype(np.int64) for i in range(nmo)]
    return (lists, X) if return_X else lists


def contract_with_TA_1st(TA, M, exch_reachable, AO_by_MO):
    """-> g (naux, sum_i |A(i)|), column k = running offset of (i, mu) (MO-major)."""
    TA, M = _f(TA), _f(M)
    nao, nmo = TA.shape
    naux = M.shape[0]
    rptr, ridx, rcol, _, _ = reachable_columns(exch_reachable)
    aptr, aidx = to_csr(AO_by_MO)
    g = np.zeros((naux, max(int(aptr[-1]), 1)), dtype=np.float64, order="F")
    lib().orc_contract_1st(c_i64(naux), c_i64(nao), c_i64(nmo), _ptr(TA), _ptr(M), _ptr(rptr), _ptr(ridx),
                           _ptr(rcol), _ptr(aptr), _ptr(aidx), _ptr(g))
    return g[:, : int(aptr[-1])]


def contract_with_TA_2nd_to_sym_dense(g, TA, AO_by_MO):
    TA, g = _f(TA), _f(g)
    nao, nmo = TA.shape
    naux = g.shape[0]
    aptr, aidx = to_csr(AO_by_MO)
    npair = nmo * (nmo + 1) // 2
    sym = np.zeros((naux, npair), dtype=np.float64, order="F")
    lib().orc_contract_2nd(c_i64(naux), c_i64(nao), c_i64(nmo), _ptr(TA), _ptr(g), _ptr(aptr), _ptr(aidx), _ptr(sym))
    return sym


def trsm_lower(L, B):
    L = _f(L)
    X = np.array(B, dtype=np.float64, order="F", copy=True)
    lib().orc_trsm_lower(c_i64(L.shape[0]), c_i64(X.shape[1]), _ptr(L), _ptr(X))
    return X


def xtx(X):
    X = _f(X)
    naux, npair = X.shape
    G = np.zeros((npair, npair), dtype=np.float64, order="F")
    lib().orc_xtx(c_i64(naux), c_i64(npair), _ptr(X), _ptr(G))
    return G


def eval_via_cholesky(sym, L):
    return xtx(trsm_lower(L, sym))


def transform_integral(M, exch_reachable, TA, S_abs, L, eps: float, AO_by_MO=None):
    """cpp:739-751 for one fragment. Returns (npair, npair) float64."""
    TA, M, L = _f(TA), _f(M), _f(L)
    nao, nmo = TA.shape
    naux = M.shape[0]
    if AO_by_MO is None:
        AO_by_MO = get_AO_per_MO(TA, S_abs, eps)
    rptr, ridx, rcol, _, _ = reachable_columns(exch_reachable)
    aptr, aidx = to_csr(AO_by_MO)
    npair = nmo * (nmo + 1) // 2
    wg = np.zeros(max(naux * int(aptr[-1]), 1), dtype=np.float64)
    ws = np.zeros(naux * npair, dtype=np.float64)
    G = np.zeros((npair, npair), dtype=np.float64, order="F")
    lib().orc_transform_integral(c_i64(naux), c_i64(nao), c_i64(nmo), _ptr(TA), _ptr(M), _ptr(rptr), _ptr(ridx),
                                 _ptr(rcol), _ptr(aptr), _ptr(aidx), _ptr(L), _ptr(wg), _ptr(ws), _ptr(G))
    return G


# --- the reference's own compiled extension -----------------------------------

def _try_import(path: Path):
    """Import in a subprocess first: a -march=native build can SIGILL on another host."""
    so = next(path.glob("eri_sparse_DF*.so"), None)
    if so is None:
        return None
    probe = subprocess.run(
        [sys.executable, "-c",
         "import importlib.util,sys,numpy as np;"
         f"s=importlib.util.spec_from_file_location('eri_sparse_DF', r'{so}');"
         "m=importlib.util.module_from_spec(s);s.loader.exec_module(m);"
         "m.get_AO_per_MO(np.eye(4),np.eye(4),0.5)"],
        capture_output=True,
    )
    if probe.returncode != 0:
        return None
    spec = importlib.util.spec_from_file_location("eri_sparse_DF", str(so))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


_REF = {}


def reference_module(prefer: str = "native"):
    """The reference's compiled extension (oracle/_ref), or None if not built / not loadable."""
    if "mod" not in _REF:
        order = ["native", "portable"] if prefer == "native" else ["portable", "native"]
        if os.environ.get("QEMB_ORACLE_REF"):
            order = [os.environ["QEMB_ORACLE_REF"]]
        mod, kind = None, None
        for k in order:
            mod = _try_import(HERE / "_ref" / k)
            if mod is not None:
                kind = k
                break
        _REF["mod"], _REF["kind"] = mod, kind
    return _REF["mod"]


def reference_kind():
    reference_module()
    return _REF.get("kind")
