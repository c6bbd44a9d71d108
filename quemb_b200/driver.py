"""All-fragment driver of the B200 sparse-DF path.

Mirrors ``_run_sparse_df_driver`` of the reference (``src/quemb/molbe/eri_sparse_DF.py:535-656``):
one-off setup (S_abs, Cholesky factor of (P|Q), screened (P|mu nu)), then one transform per fragment
whose ``(npair, npair)`` result goes to the caller's sink (``file_eri_handler.create_dataset`` in
QuEmb).  Differences, all behind the same seam:

* (P|mu nu), L and S_abs are uploaded once and stay in HBM; per fragment only ``TA`` goes up and G comes
  back (the reference's GPU variant re-uploads (P|ij) and mallocs per fragment, ``cpp:642-660``);
* the metric is applied once to (P|mu nu) (``whiten=True``, design option O1) instead of once per
  fragment;
* calls are pipelined: while the sink consumes fragment f-1 and the copy engine drains G(f), the
  compute stream already works on fragment f+1;
* fragments are partitioned over ranks (one process per GPU, ``torch.distributed``) or over several
  devices of one process (one thread per device, dynamic queue) with no data-path collective.
"""

from __future__ import annotations

import threading
from typing import Callable, Sequence

import numpy as np

from . import _lib
from . import eri_sparse_DF as q

Sink = Callable[[int, np.ndarray], None]


def fragment_cost(naux: int, nmo: int, nao: int | None = None) -> float:
    """Relative cost model used for partitioning: K3 dominates (naux*npair*(npair+1))."""
    npair = nmo * (nmo + 1) // 2
    return float(naux) * npair * (npair + 1)


def partition_lpt(costs: Sequence[float], n_parts: int) -> list[list[int]]:
    """Longest-processing-time greedy partition of fragment indices over ``n_parts`` GPUs."""
    order = sorted(range(len(costs)), key=lambda i: (-costs[i], i))
    loads = [0.0] * n_parts
    parts: list[list[int]] = [[] for _ in range(n_parts)]
    for i in order:
        k = min(range(n_parts), key=lambda j: (loads[j], j))
        parts[k].append(i)
        loads[k] += costs[i]
    return [sorted(p) for p in parts]


class FragmentERIBuilder:
    """Device-resident state of one molecule on ONE GPU + the pipelined per-fragment loop."""

    def __init__(self, device: int = 0, max_nmo: int | None = None, n_host_buffers: int = 3):
        self.ctx = q.default_context(device)
        self.ws = q.Workspace(self.ctx)
        self.tensor: q.SemiSparseSym3DTensor | None = None
        self.metric: q.GPU_MatrixHandle | None = None
        self.screen = None
        self._bufs: list[np.ndarray] = []
        self._n_host_buffers = n_host_buffers
        if max_nmo:
            self.reserve_host(max_nmo)

    # -- one-off staging (timed separately by the benchmark, like libcint in the north star) --------
    def stage(self, unique_dense_data: np.ndarray | None, shape, exch_reachable, L_PQ, S_abs, *, whiten: bool = True,
              tensor: q.SemiSparseSym3DTensor | None = None) -> None:
        if tensor is None:
            tensor = q.SemiSparseSym3DTensor(unique_dense_data, shape, exch_reachable, ctx=self.ctx)
        self.tensor = tensor
        self.metric = q.GPU_MatrixHandle(L_PQ, ctx=self.ctx) if not isinstance(L_PQ, q.GPU_MatrixHandle) else L_PQ
        self.screen = q._ScreenHandle(np.asarray(S_abs), self.ctx) if not isinstance(S_abs, q._ScreenHandle) else S_abs
        if whiten and not self.tensor.whitened:
            self.tensor.whiten(self.metric)

    def reserve_host(self, max_nmo: int) -> None:
        npair = max_nmo * (max_nmo + 1) // 2
        need = npair * npair
        if self._bufs and self._bufs[0].size >= need:
            return
        for b in self._bufs:
            _lib.pinned_free(b)
        self._bufs = [_lib.pinned_empty((need,)) for _ in range(self._n_host_buffers)]

    def close(self) -> None:
        self.ctx.synchronize()
        for b in self._bufs:
            _lib.pinned_free(b)
        self._bufs = []

    # -- the loop -------------------------------------------------------------------------------------
    def build(self, TAs: Sequence[np.ndarray], MO_coeff_epsilon: float, sink: Sink | None = None,
              indices: Sequence[int] | None = None) -> None:
        """Transform the fragments ``indices`` (default: all) and hand every G to ``sink(i, G)``.
        ``G`` is a view of a pinned staging buffer, valid only during the sink call."""
        assert self.tensor is not None, "stage() first"
        idx = list(range(len(TAs))) if indices is None else list(indices)
        if not idx:
            return
        self.reserve_host(max(TAs[i].shape[1] for i in idx))
        nb = len(self._bufs)
        metric = None if self.tensor.whitened else self.metric
        pending: list[tuple[int, int, np.ndarray]] = []  # (fragment, ticket, view)

        def drain_one():
            i, ticket, view = pending.pop(0)
            self.ws.wait_output(ticket)
            if sink is not None:
                sink(i, view)

        for k, i in enumerate(idx):
            TA = TAs[i]
            nmo = TA.shape[1]
            npair = nmo * (nmo + 1) // 2
            if len(pending) >= nb:  # the buffer about to be reused must have been consumed
                drain_one()
            view = self._bufs[k % nb][: npair * npair].reshape(npair, npair)
            q.transform_integral(self.tensor, TA, self.screen, metric, MO_coeff_epsilon, ws=self.ws, out=view,
                                 synchronize=False)
            pending.append((i, self.ws.last_ticket, view))
            if len(pending) >= nb:
                drain_one()
        while pending:
            drain_one()

    def build_device(self, TAs: Sequence[np.ndarray], MO_coeff_epsilon: float, out_dev_ptr: int, ld_out: int | None = None,
                     indices: Sequence[int] | None = None) -> None:
        """Same loop with G left in caller-owned device memory (every fragment overwrites ``out``):
        the kernel-only path the benchmark's ``value`` is measured on."""
        assert self.tensor is not None, "stage() first"
        idx = list(range(len(TAs))) if indices is None else list(indices)
        metric = None if self.tensor.whitened else self.metric
        for i in idx:
            nmo = TAs[i].shape[1]
            npair = nmo * (nmo + 1) // 2
            q.transform_integral_device(self.tensor, TAs[i], self.screen, metric, MO_coeff_epsilon, out_dev_ptr,
                                        ld_out or npair, ws=self.ws)


def build_all_fragments(unique_dense_data, shape, exch_reachable, L_PQ, S_abs, TAs, MO_coeff_epsilon, sink: Sink,
                        *, devices: Sequence[int] = (0,), whiten: bool = True) -> None:
    """One process, one thread per device, fragments handed out dynamically (largest first)."""
    naux = shape[0]
    order = sorted(range(len(TAs)), key=lambda i: -fragment_cost(naux, TAs[i].shape[1]))
    lock = threading.Lock()
    errors: list[BaseException] = []

    def worker(dev: int):
        try:
            b = FragmentERIBuilder(dev)
            b.stage(unique_dense_data, shape, exch_reachable, L_PQ, S_abs, whiten=whiten)
            while True:
                with lock:
                    if not order:
                        break
                    i = order.pop(0)
                b.build(TAs, MO_coeff_epsilon, sink, indices=[i])
            b.close()
        except BaseException as e:  # noqa: BLE001
            errors.append(e)

    threads = [threading.Thread(target=worker, args=(d,)) for d in devices]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    if errors:
        raise errors[0]


# ---------------------------------------------------------------------------
# QuEmb seam (needs quemb + pyscf + h5py in the environment)
# ---------------------------------------------------------------------------


def transform_sparse_DF_integral_b200(mf, Fobjs, auxbasis, file_eri_handler, AO_coeff_epsilon: float,
                                      MO_coeff_epsilon: float, n_threads: int = 1, precompute_P_mu_nu: bool = True,
                                      devices: Sequence[int] = (0,)) -> None:
    """Drop-in sibling of ``transform_sparse_DF_integral_gpu`` (``eri_sparse_DF.py:683-704``).

    Host-side one-offs stay exactly the reference's own functions (``approx_S_abs``, libcint
    ``int2c2e`` / ``int3c2e`` through ``get_sparse_P_mu_nu``, SciPy Cholesky); everything after them
    runs on the B200.  Writes dataset ``fobj.dname`` of shape ``(npair, npair)`` float64 per fragment
    (``eri_sparse_DF.py:630-631``)."""
    from pyscf.ao2mo.addons import restore  # noqa: PLC0415
    from pyscf.df.addons import make_auxmol  # noqa: PLC0415
    from quemb.molbe import eri_sparse_DF as ref  # noqa: PLC0415
    from scipy.linalg import cholesky  # noqa: PLC0415

    mol = mf.mol
    auxmol = make_auxmol(mol, auxbasis=auxbasis)
    S_abs = ref.approx_S_abs(mol)
    PQ = auxmol.intor("int2c2e")
    L = cholesky(PQ, lower=True)

    def sink_for(fobjs):
        def sink(i: int, G: np.ndarray) -> None:
            eri = restore("4", G, fobjs[i].TA.shape[1])
            file_eri_handler.create_dataset(fobjs[i].dname, data=eri)
        return sink

    lock = threading.Lock()

    def locked(sink):
        def s(i, G):
            with lock:  # h5py serialises writers anyway
                sink(i, G)
        return s

    TAs = [np.asfortranarray(f.TA) for f in Fobjs]
    if precompute_P_mu_nu:
        exch = ref._get_AO_per_AO(S_abs, AO_coeff_epsilon, None)
        exch_lists = [np.asarray(exch[mu], dtype=np.int64) for mu in range(mol.nao)]
        # the reference's own libcint driver fills the host staging view of our tensor
        P_ref = ref.get_sparse_P_mu_nu(mol, auxmol, exch)
        build_all_fragments(P_ref.unique_dense_data, (auxmol.nao, mol.nao, mol.nao), exch_lists, L, S_abs, TAs,
                            MO_coeff_epsilon, locked(sink_for(Fobjs)), devices=devices, whiten=True)
    else:
        b = FragmentERIBuilder(devices[0])
        for i, fobj in enumerate(Fobjs):
            exch = ref._get_AO_per_AO(S_abs, AO_coeff_epsilon, fobj.TA)
            P_ref = ref.get_sparse_P_mu_nu(mol, auxmol, exch)
            exch_lists = [np.asarray(exch[mu], dtype=np.int64) for mu in range(mol.nao)]
            b.stage(P_ref.unique_dense_data, (auxmol.nao, mol.nao, mol.nao), exch_lists, L, S_abs, whiten=False)
            b.build(TAs, MO_coeff_epsilon, sink_for(Fobjs), indices=[i])
        b.close()
