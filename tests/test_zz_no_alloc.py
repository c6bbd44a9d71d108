"""GPU: ``qemb_gram_noalloc`` -- the K3 entry point that allocates nothing (split-K scratch in a static ``__device__``
array, ``csrc/gram_noalloc.cu``; no counterpart in the reference) -- against the existing K3 entry point
(``qemb_fragment_transform`` / ``qemb_fragment_gram``, workspace scratch from cudaMalloc) on the same X, through the C ABI.

Collected last on purpose (a failure here must not hide the other tests); passes alone in a fresh
process (the first call does the per-device set-up) and after the rest of the suite.
"""
from __future__ import annotations

import ctypes

import numpy as np
import pytest

from quemb_b200 import _lib, synthetic as syn

from helpers import Golden

pytestmark = pytest.mark.gpu

TOL = 1e-10  # the suite's max-abs bar on (ij|kl)
c_i64, c_p = ctypes.c_int64, ctypes.c_void_p


@pytest.fixture(scope="module")
def q():
    from quemb_b200 import eri_sparse_DF as mod

    mod.default_context(0)  # raises QembNoDevice without a B200: no CPU fallback
    return mod


def _existing_X_and_G(q, T, TA, S_abs, L, eps, naux):
    """X = L^-1 (P|ij) and G = X^T X of one fragment from the EXISTING entry points (workspace path)."""
    ws = q.transform_integral_resident(T, TA, S_abs, L, eps)
    nmo = TA.shape[1]
    npair = nmo * (nmo + 1) // 2
    X = np.empty((naux, npair), order="F")
    _lib.check(_lib.lib().qemb_fragment_get_sym(ws._h, _lib.ptr(X), c_i64(naux)))
    return X, ws.G(npair)


def _noalloc_gram(torch, X: np.ndarray, ld_extra: int = 0):
    """G from qemb_gram_noalloc; X travels with NaN in the rows [naux, ldx) and G with a sentinel in its padding rows:
    neither may be touched."""
    naux, npair = X.shape
    ldx = naux + (naux % 2) + 2 * ld_extra
    ldg = npair + 3
    Xh = np.full((npair, ldx), np.nan)  # C order (npair, ldx) == column-major ldx x npair
    Xh[:, :naux] = X.T
    Xd = torch.from_numpy(Xh).to("cuda:0")
    Gd = torch.full((npair, ldg), -7.0, dtype=torch.float64, device="cuda:0")
    stream = torch.cuda.current_stream(0).cuda_stream
    with torch.cuda.device(0):
        _lib.check(_lib.lib().qemb_gram_noalloc(c_p(Xd.data_ptr()), c_i64(ldx), c_i64(naux), c_i64(npair),
                                               c_p(Gd.data_ptr()), c_i64(ldg), c_p(stream)))
        torch.cuda.synchronize(0)
    Gh = Gd.cpu().numpy()  # Gh[c, r] = G[r + ldg * c]
    assert np.all(Gh[:, npair:] == -7.0), "rows past npair of the output were written"
    return Gh[:, :npair].T.copy()


def _check(G_new, G_old, what):
    assert np.isfinite(G_new).all(), f"{what}: non-finite entries (padding rows of X read, or stale scratch)"
    dev = float(np.abs(G_new - G_old).max())
    assert dev <= TOL, f"{what}: max |G_noalloc - G_existing| = {dev:.3e} > {TOL}"
    assert np.array_equal(G_new, G_new.T), f"{what}: not symmetric"
    return dev


def test_gram_noalloc_matches_existing_entry_point_on_consecutive_calls(q):
    torch = pytest.importorskip("torch")
    max_npair, scratch = c_i64(), c_i64()
    _lib.check(_lib.lib().qemb_gram_noalloc_limits(ctypes.byref(max_npair), ctypes.byref(scratch)))
    assert (max_npair.value, scratch.value) == (2048, 32 << 20)

    p = syn.make_chain_problem(n_units=8, ao_per_unit=8, n_end=1, naux=203, frag_halfwidth=1, n_bath=16, sigma=0.5,
                               seed=11)
    ctx = q.default_context(0)
    T = q.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable, ctx=ctx)
    L = q.GPU_MatrixHandle(p.L_PQ, ctx=ctx)
    # two fragments of different size (npair 861: 28 tiles, split-K over the scratch array; npair 561) ...
    cases = []
    for f in (1, 0):
        X, G = _existing_X_and_G(q, T, p.TAs[f], p.S_abs, L, p.MO_coeff_epsilon, p.naux)
        cases.append((f"chain fragment {f} (npair {X.shape[1]}, naux {X.shape[0]})", X, G))
    # ... and a tiny golden one (npair 28, naux 24: one tile, k tail inside the first 16-row step's neighbour)
    g = Golden("asym_sources")
    Tg = q.SemiSparseSym3DTensor(g.M, (g.naux, g.nao, g.nao), [x.tolist() for x in g.exch])
    X, G = _existing_X_and_G(q, Tg, g.frag(0)["TA"], g.S_abs, g.L, g.eps, g.naux)
    assert np.abs(G - g.frag(0)["G"]).max() <= TOL  # the existing path itself against the reference's golden
    cases.append((f"golden asym_sources (npair {X.shape[1]}, naux {X.shape[0]})", X, G))

    # back-to-back calls that all share the one scratch array: large, small, tiny, then the large one again over the
    # leftovers of the smaller ones
    worst = 0.0
    for k, (what, X, G) in enumerate(cases + cases[:1]):
        worst = max(worst, _check(_noalloc_gram(torch, X, ld_extra=k % 2), G, f"call {k}: {what}"))
    print(f"qemb_gram_noalloc: 4 consecutive calls, max |G_noalloc - G_existing| = {worst:.3e}")


def test_gram_noalloc_rejects_sizes_beyond_the_scratch_array(q):
    torch = pytest.importorskip("torch")
    buf = torch.zeros(64, dtype=torch.float64, device="cuda:0")
    L = _lib.lib()
    rc = L.qemb_gram_noalloc(c_p(buf.data_ptr()), c_i64(16), c_i64(16), c_i64(2049), c_p(buf.data_ptr()), c_i64(2049),
                             c_p(None))
    assert rc == _lib.QEMB_EINVAL
    msg = _lib.last_error()
    assert "2049" in msg and "maximum of 2048" in msg and "32 MiB" in msg
    with pytest.raises(_lib.QembError, match="ldx"):
        _lib.check(L.qemb_gram_noalloc(c_p(buf.data_ptr()), c_i64(15), c_i64(16), c_i64(2), c_p(buf.data_ptr()),
                                       c_i64(2), c_p(None)))
    with pytest.raises(_lib.QembError, match="null"):
        _lib.check(L.qemb_gram_noalloc(c_p(None), c_i64(16), c_i64(16), c_i64(2), c_p(buf.data_ptr()), c_i64(2),
                                       c_p(None)))
