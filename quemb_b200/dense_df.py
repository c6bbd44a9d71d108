"""``int-direct-DF`` on the sparse-DF kernels: the unscreened limit R(mu) = A(i) = everything.

Reference: ``integral_direct_DF`` (``src/quemb/molbe/eri_onthefly.py:45-145``): (ij|P) = TA^T (mu nu|P) TA for ALL AO
pairs, ``solve_triangular(low, .)``, ``bb.T @ bb``, ``restore("4", ...)``.  With full lists the screened path
computes exactly that, so the same K1 / K2a / K2b / K3 kernels serve it and give an independent cross-check of
the sparse path on small molecules (``tests/eri_onthefly_test.py:64-68`` of the reference pins its energy).
"""

from __future__ import annotations

import numpy as np

from . import eri_sparse_DF as q


def pack_dense(pqL: np.ndarray) -> tuple[np.ndarray, list[np.ndarray]]:
    """(naux, nao, nao) symmetric-in-(mu,nu) dense tensor -> unique_dense_data (naux, nao(nao+1)/2) in the
    reference's column order (mu ascending, nu <= mu; ``eri_sparse_DF.cpp:248-258``) + the full lists."""
    naux, nao, nao2 = pqL.shape
    assert nao == nao2
    mu, nu = np.tril_indices(nao)  # row-major over the lower triangle: mu ascending, then nu = 0..mu
    data = np.asfortranarray(pqL[:, mu, nu])
    full = [np.arange(nao, dtype=np.int64) for _ in range(nao)]
    return data, full


class DenseDF:
    """Device-resident dense (P|mu nu) + metric; ``transform(TA)`` returns the s4 ``(npair, npair)`` ERIs that
    ``integral_direct_DF`` writes to ``file_eri`` (``eri_onthefly.py:139-145``)."""

    def __init__(self, pqL: np.ndarray, low: np.ndarray, ctx: q.Context | None = None):
        naux, nao, _ = pqL.shape
        data, full = pack_dense(np.asarray(pqL, dtype=np.float64))
        self.ctx = ctx or q.default_context()
        self.tensor = q.SemiSparseSym3DTensor(data, (naux, nao, nao), full, ctx=self.ctx)
        self.tensor.whiten(low)
        self._ones = q._ScreenHandle(np.ones((nao, nao)), self.ctx)  # with eps = 0 nothing is screened

    def transform(self, TA: np.ndarray, out: np.ndarray | None = None) -> np.ndarray:
        return q.transform_integral(self.tensor, TA, self._ones, None, 0.0, out=out)


def integral_direct_DF(pqL: np.ndarray, low: np.ndarray, TAs, sink) -> None:
    """All fragments: ``sink(i, eri)`` takes the place of ``file_eri.create_dataset(Fobjs[i].dname, data=eri)``."""
    d = DenseDF(pqL, low)
    for i, TA in enumerate(TAs):
        sink(i, d.transform(np.asfortranarray(TA)))
