"""CPU: the C-ABI library loads, exports every symbol include/qemb_b200.h declares, its host-side
index maps are bit-exact, and the compute entry points fail loudly without a GPU (no CPU fallback)."""
from __future__ import annotations

import ctypes

import numpy as np
import pytest

from quemb_b200 import _lib, indexing

from helpers import CASES, Golden


@pytest.fixture(scope="module")
def L():
    if not _lib.LIB_PATH.exists():
        _lib.build()
    return _lib.lib()


def test_exports_every_declared_symbol(L):
    syms = _lib.declared_symbols()
    assert len(syms) >= 40
    missing = [s for s in syms if not hasattr(L, s)]
    assert not missing
    assert set(syms) == set(_lib._SIGS), "ctypes table and header disagree"


def test_index_maps_host_side(L):
    for a in range(40):
        for b in range(40):
            assert L.qemb_ravel_symmetric(a, b) == indexing.ravel_symmetric(a, b)
    lo, hi = ctypes.c_int64(), ctypes.c_int64()
    for i in list(range(3000)) + [2**40 + 12345, 2**52 - 1]:
        L.qemb_unravel_symmetric(i, ctypes.byref(lo), ctypes.byref(hi))
        assert (lo.value, hi.value) == indexing.unravel_symmetric(i)


@pytest.mark.parametrize("name", CASES)
def test_extract_unique_matches_reference_golden(L, name):
    g = Golden(name)
    eptr, eidx = indexing.to_csr(g.exch)
    uptr = np.zeros(g.nao + 1, dtype=np.int64)
    n = L.qemb_extract_unique(g.nao, _lib.ptr(eptr), _lib.ptr(eidx), _lib.ptr(uptr), None)
    uidx = np.zeros(max(n, 1), dtype=np.int64)
    L.qemb_extract_unique(g.nao, _lib.ptr(eptr), _lib.ptr(eidx), _lib.ptr(uptr), _lib.ptr(uidx))
    assert indexing.from_csr(uptr, uidx[:n]) == [u.tolist() for u in g.uniq]


def test_no_cpu_fallback(L):
    if _lib.device_count() > 0:
        pytest.skip("a GPU is visible")
    h = ctypes.c_void_p()
    rc = L.qemb_ctx_create(0, ctypes.byref(h))
    assert rc == _lib.QEMB_ENODEV
    assert b"no CUDA device" in L.qemb_last_error() or b"cudaGetDeviceCount" in L.qemb_last_error()
    from quemb_b200 import eri_sparse_DF as q

    with pytest.raises(_lib.QembNoDevice):
        q.get_AO_per_MO(np.eye(4), np.eye(4), 1e-5)


@pytest.mark.parametrize("n", [1, 7, 255, 1030, 2051])
def test_host_mirror_lower(n):
    """Host half of output mode 1 (no GPU needed): upper triangle rebuilt from the lower one, bit-exact."""
    import ctypes

    L = _lib.lib()
    rng = np.random.default_rng(n)
    A = rng.standard_normal((n, n))
    S = np.tril(A) + np.tril(A, -1).T
    ld = n + 3
    buf = np.full((n, ld), np.nan)          # column c of the column-major matrix = buf[c, :n]
    buf[:, :n] = np.where(np.arange(n)[None, :] >= np.arange(n)[:, None], S, np.nan)  # rows r >= c only
    _lib.check(L.qemb_host_mirror_lower(_lib.ptr(buf), ctypes.c_int64(ld), ctypes.c_int64(n)))
    np.testing.assert_array_equal(buf[:, :n], S)
    nt = ctypes.c_int()
    _lib.check(L.qemb_host_threads(ctypes.byref(nt)))
    assert nt.value >= 1


def test_host_mirror_lower_avx2_path_in_a_fresh_process():
    """The AVX2 mirror (hosts without AVX-512) is selected once per process: exercise it through QEMB_NO_AVX512."""
    import os
    import subprocess
    import sys

    code = (
        "import sys, ctypes, numpy as np; sys.path.insert(0, %r);"
        "from quemb_b200 import _lib; L = _lib.lib();"
        "rng = np.random.default_rng(0)\n"
        "for n in (7, 300, 1030, 2051):\n"
        "    G = np.asfortranarray(rng.standard_normal((n, n))); want = np.tril(G) + np.tril(G, -1).T\n"
        "    assert L.qemb_host_mirror_lower(ctypes.c_void_p(G.ctypes.data), ctypes.c_int64(n), ctypes.c_int64(n)) == 0\n"
        "    assert np.array_equal(G, want), n\n"
        "print('ok')" % str(_lib.ROOT))
    r = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, QEMB_NO_AVX512="1"), capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.strip() == "ok", r.stderr


def test_log_level_default_and_roundtrip_in_a_fresh_process():
    """``get_log_level`` / ``set_log_level`` (``indexers.hpp:34-46``): process global, default Error = 40."""
    import subprocess
    import sys

    code = ("import sys; sys.path.insert(0, %r); from quemb_b200 import eri_sparse_DF as q;"
            "a = q.get_log_level(); q.set_log_level(20); b = q.get_log_level(); q.set_log_level(40); print(a, b, q.get_log_level())"
            % str(_lib.ROOT))
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.split() == ["40", "20", "40"], r.stderr


def test_gram_noalloc_argument_checks_need_no_device(L):
    """``qemb_gram_noalloc`` rejects bad arguments -- above all sizes beyond its fixed scratch array -- before it touches
    CUDA, and on a box without a GPU a well-formed call fails loudly instead of falling back to anything."""
    i64, vp = ctypes.c_int64, ctypes.c_void_p
    mx, sb = i64(), i64()
    assert L.qemb_gram_noalloc_limits(ctypes.byref(mx), ctypes.byref(sb)) == 0
    assert mx.value == 2048 and sb.value == 32 << 20 and 8 * mx.value ** 2 <= sb.value
    fake = vp(4096)  # never dereferenced on these paths
    assert L.qemb_gram_noalloc(vp(None), i64(16), i64(16), i64(8), fake, i64(8), vp(None)) == _lib.QEMB_EINVAL
    assert b"null" in L.qemb_last_error()
    assert L.qemb_gram_noalloc(fake, i64(16), i64(16), i64(mx.value + 1), fake, i64(mx.value + 1), vp(None)) == _lib.QEMB_EINVAL
    msg = L.qemb_last_error().decode()
    assert "exceeds the maximum of 2048" in msg and "32 MiB" in msg
    assert L.qemb_gram_noalloc(fake, i64(15), i64(16), i64(8), fake, i64(8), vp(None)) == _lib.QEMB_EINVAL   # ldx < naux
    assert L.qemb_gram_noalloc(fake, i64(17), i64(16), i64(8), fake, i64(8), vp(None)) == _lib.QEMB_EINVAL   # odd ldx
    assert L.qemb_gram_noalloc(fake, i64(16), i64(16), i64(8), fake, i64(7), vp(None)) == _lib.QEMB_EINVAL   # ldg < npair
    assert L.qemb_gram_noalloc(fake, i64(16), i64(0), i64(8), fake, i64(8), vp(None)) == _lib.QEMB_EINVAL    # naux < 1
    if _lib.device_count() == 0:
        rc = L.qemb_gram_noalloc(fake, i64(16), i64(16), i64(8), fake, i64(8), vp(None))
        assert rc in (_lib.QEMB_ENODEV, _lib.QEMB_ECUDA) and L.qemb_last_error()
