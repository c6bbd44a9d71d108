"""All-fragment driver of the B200 sparse-DF path.

Mirrors ``_run_sparse_df_driver`` of the reference (``src/quemb/molbe/eri_sparse_DF.py:535-656``):
one-off setup (S_abs, Cholesky factor of (P|Q), screened (P|mu nu)), then one transform per fragment
whose ``(npair, npair)`` result goes to the caller's sink (``file_eri_handler.create_dataset`` in
QuEmb).  Differences, all behind the same seam:

* (P|mu nu), L and S_abs are uploaded once and stay in HBM; per fragment only ``TA`` goes up and G comes
  back (the reference's GPU variant re-uploads (P|ij) and mallocs per fragment, ``cpp:642-660``);
* (P|mu nu) is staged block by block straight into HBM (``staging.get_sparse_P_mu_nu``), never through a host
  copy and never through the reference's per-pair Python scatter (``eri_sparse_DF.py:476-488``);
* the metric is applied once to (P|mu nu) (``whiten=True``, design option O1) instead of once per
  fragment;
* calls are pipelined: while the sink consumes fragment f-1 and the copy engine drains G(f), the
  compute stream already works on fragment f+1; a sink may return a ``Future`` (the background dataset writer
  does) and the pinned buffer it was handed is only reused once that future is done;
* fragments are partitioned over ranks (one process per GPU, ``torch.distributed``) or over several
  devices of one process (one thread per device, dynamic queue) with no data-path collective;
* a consumer that only needs J/K-type contractions of every fragment's (ij|kl) (``BE.initialize``,
  ``mbe.py:1146-1170``) runs where the matrix lives (:meth:`FragmentERIBuilder.build_jk`): only ``nmo x nmo``
  matrices leave the device.
"""

from __future__ import annotations

import logging
import threading
from concurrent.futures import Future
from typing import Callable, Sequence

import numpy as np

from . import _lib
from . import eri_sparse_DF as q
from . import staging

logger = logging.getLogger(__name__)

Sink = Callable[[int, np.ndarray], "Future | None"]


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


class _Lane:
    """One caller's worth of per-fragment state: a workspace (own CUDA streams) + its pinned output ring."""

    def __init__(self, ctx: q.Context, ws: q.Workspace | None = None):
        self.ws = ws or q.Workspace(ctx)
        self.bufs: list[np.ndarray] = []

    def reserve(self, need: int, n: int) -> None:
        if self.bufs and self.bufs[0].size >= need and len(self.bufs) == n:
            return
        self.free()
        self.bufs = [_lib.pinned_empty((need,)) for _ in range(n)]

    def free(self) -> None:
        for b in self.bufs:
            _lib.pinned_free(b)
        self.bufs = []


class FragmentERIBuilder:
    """Device-resident state of one molecule on ONE GPU + the pipelined per-fragment loop."""

    def __init__(self, device: int = 0, max_nmo: int | None = None, n_host_buffers: int = 3):
        self.ctx = q.default_context(device)
        self.ws = q.Workspace(self.ctx)
        self._lanes: list[_Lane] = [_Lane(self.ctx, self.ws)]
        self._jk_pool: np.ndarray | None = None
        self.tensor: q.SemiSparseSym3DTensor | None = None
        self.metric: q.GPU_MatrixHandle | None = None
        self.screen = None
        # the library keeps completion state for `output_ring_depth()` pending host outputs per workspace; more
        # buffers than that would let a ticket be recycled before anybody waited for it
        depth = q.output_ring_depth()
        if n_host_buffers < 1:
            raise ValueError("n_host_buffers must be >= 1")
        if n_host_buffers > depth:
            logger.warning("n_host_buffers=%d clamped to the library's output ring depth %d", n_host_buffers, depth)
        self._n_host_buffers = min(n_host_buffers, depth)
        self.borderline_total = 0
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

    def stage_distributed(self, unique_dense_data: np.ndarray, L_PQ, S_abs, dist, *, tensor: q.SemiSparseSym3DTensor,
                          group=None) -> None:
        """Fragment-parallel ranks, every one holding the host array: rank r uploads only ITS 1/world share of the
        columns of (P|mu nu), whitens that share, and the shares are exchanged on the device with NCCL broadcasts
        over NVLink (``SemiSparseSym3DTensor.whiten_distributed``) -- one host->device pass of the tensor per NODE
        instead of one per rank, and 1/world of the metric solve each."""
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        n = tensor._n_unique
        c0, c1 = (n * rank) // world, (n * (rank + 1)) // world
        self.tensor = tensor
        self.metric = q.GPU_MatrixHandle(L_PQ, ctx=self.ctx) if not isinstance(L_PQ, q.GPU_MatrixHandle) else L_PQ
        self.screen = q._ScreenHandle(np.asarray(S_abs), self.ctx) if not isinstance(S_abs, q._ScreenHandle) else S_abs
        tensor.set_whitened(False)
        if c1 > c0:
            tensor.upload(unique_dense_data[:, c0:c1], col0=c0)
        tensor.whiten_distributed(self.metric, dist, group=group)

    @property
    def _bufs(self) -> list[np.ndarray]:
        return self._lanes[0].bufs

    def tune_output(self) -> dict[str, float]:
        """Let the library balance PCIe against the host mirror threads for this rank (``Context.autotune_output``);
        call on all ranks of a node at the same moment."""
        return self.ctx.autotune_output()

    def build_dynamic(self, TAs: Sequence[np.ndarray], MO_coeff_epsilon: float, sink: Sink | None, queue,
                      n_threads: int = 1) -> list[int]:
        """:meth:`build` over a queue shared with other builders (:class:`DistributedQueue` across ranks): returns
        the fragments this builder ended up transforming."""
        assert self.tensor is not None, "stage() first"
        self._build_from(queue, max(t.shape[1] for t in TAs), TAs, MO_coeff_epsilon, sink, n_threads)
        return list(getattr(queue, "taken", []))

    def reserve_host(self, max_nmo: int, lanes: int = 1) -> None:
        npair = max_nmo * (max_nmo + 1) // 2
        while len(self._lanes) < lanes:
            self._lanes.append(_Lane(self.ctx))
        for lane in self._lanes[:lanes]:
            lane.reserve(npair * npair, self._n_host_buffers)

    def close(self) -> None:
        self.ctx.synchronize()
        for lane in self._lanes:
            lane.free()
        if self._jk_pool is not None:
            _lib.pinned_free(self._jk_pool)
            self._jk_pool = None

    # -- the loop -------------------------------------------------------------------------------------
    def build(self, TAs: Sequence[np.ndarray], MO_coeff_epsilon: float, sink: Sink | None = None,
              indices: Sequence[int] | None = None, n_threads: int = 1) -> None:
        """Transform the fragments ``indices`` (default: all) and hand every G to ``sink(i, G)``.
        ``G`` is a view of a pinned staging buffer, valid during the sink call -- or, when the sink returns a
        ``concurrent.futures.Future`` (e.g. :class:`quemb_b200.eri_file.AsyncDatasetWriter`), until that future
        is done: the buffer is not handed to the device again before.

        ``n_threads > 1`` is the reference's ``ThreadPoolExecutor(n_threads)`` around its per-fragment worker
        (``eri_sparse_DF.py:649-653``): that many host threads, each with a workspace (= CUDA streams) and an output
        ring of its own, take fragments from a shared queue; the library calls release the GIL, the kernels of
        different lanes overlap on the device -- what keeps 148 SMs busy when single fragments are too small to
        (nmo <~ 100).  ``sink`` calls are serialised."""
        assert self.tensor is not None, "stage() first"
        idx = list(range(len(TAs))) if indices is None else list(indices)
        if not idx:
            return
        n_threads = max(1, min(int(n_threads), len(idx)))
        self._build_from(iter(idx), max(TAs[i].shape[1] for i in idx), TAs, MO_coeff_epsilon, sink, n_threads)

    def _build_from(self, it, max_nmo: int, TAs, MO_coeff_epsilon: float, sink: Sink | None, n_threads: int = 1) -> None:
        """The fragment loop over an ITERATOR of fragment indices (possibly shared with other devices' builders and
        then already thread-safe): every lane keeps its own pipeline full until the iterator is exhausted."""
        self.reserve_host(max_nmo, lanes=n_threads)
        if n_threads == 1:
            self._build_lane(self._lanes[0], it, TAs, MO_coeff_epsilon, sink)
            return
        it_lock, sink_lock = threading.Lock(), threading.Lock()

        class Shared:
            def __iter__(self):
                return self

            def __next__(self):
                with it_lock:
                    return next(it)

        def locked_sink(i, G):
            with sink_lock:
                return sink(i, G)

        errors: list[BaseException] = []

        def run(lane):
            try:
                self._build_lane(lane, Shared(), TAs, MO_coeff_epsilon, None if sink is None else locked_sink)
            except BaseException as e:  # noqa: BLE001
                errors.append(e)

        threads = [threading.Thread(target=run, args=(lane,)) for lane in self._lanes[:n_threads]]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        if errors:
            raise errors[0]

    def _build_lane(self, lane: _Lane, fragments, TAs, MO_coeff_epsilon: float, sink: Sink | None) -> None:
        nb = len(lane.bufs)
        ws = lane.ws
        metric = None if self.tensor.whitened else self.metric
        pending: list[tuple[int, int, np.ndarray, int]] = []  # (fragment, ticket, view, buffer)
        busy: list[Future | None] = [None] * nb  # sink still reading buffer k

        def drain_one():
            i, ticket, view, kb = pending.pop(0)
            ws.wait_output(ticket)
            if sink is not None:
                fut = sink(i, view)
                busy[kb] = fut if isinstance(fut, Future) else None

        for k, i in enumerate(fragments):
            TA = TAs[i]
            nmo = TA.shape[1]
            npair = nmo * (nmo + 1) // 2
            kb = k % nb
            if len(pending) >= nb:  # the buffer about to be reused must have been consumed
                drain_one()
            if busy[kb] is not None:
                busy[kb].result()
                busy[kb] = None
            view = lane.bufs[kb][: npair * npair].reshape(npair, npair)
            q.transform_integral(self.tensor, TA, self.screen, metric, MO_coeff_epsilon, ws=ws, out=view,
                                 synchronize=False)
            self._note_borderline(i, ws)
            pending.append((i, ws.last_ticket, view, kb))
            if len(pending) >= nb:
                drain_one()
        while pending:
            drain_one()
        for f in busy:
            if f is not None:
                f.result()

    def _note_borderline(self, i: int, ws: q.Workspace | None = None) -> None:
        """Screening entries within 1e-12 of the threshold may fall on the other side in an implementation that sums
        in another order (the reference's Eigen GEMM is blocked, FMA and -ffast-math): say so instead of silently
        producing lists that can differ from the reference's."""
        n = (ws or self.ws).borderline
        if n:
            self.borderline_total += n
            logger.warning("fragment %d: %d screening entries within 1e-12 of MO_coeff_epsilon; A(i) may differ from "
                           "an implementation that sums S_abs*TA in another order", i, n)

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

    def build_resident(self, TAs: Sequence[np.ndarray], MO_coeff_epsilon: float, indices: Sequence[int] | None = None,
                       n_threads: int = 1) -> None:
        """Transforms with (ij|kl) left in each lane's workspace and nothing delivered: the kernel-only path for
        BATCHES of small fragments (``n_threads`` lanes overlap on the device).  Asynchronous per lane; callers
        bracket it with ``ctx.mark`` / ``ctx.synchronize``."""
        assert self.tensor is not None, "stage() first"
        idx = list(range(len(TAs))) if indices is None else list(indices)
        n_threads = max(1, min(int(n_threads), max(len(idx), 1)))
        while len(self._lanes) < n_threads:
            self._lanes.append(_Lane(self.ctx))
        metric = None if self.tensor.whitened else self.metric
        it, lock = iter(idx), threading.Lock()
        errors: list[BaseException] = []

        def run(lane):
            try:
                while True:
                    with lock:
                        i = next(it, None)
                    if i is None:
                        return
                    q.transform_integral_resident(self.tensor, TAs[i], self.screen, metric, MO_coeff_epsilon, ws=lane.ws)
            except BaseException as e:  # noqa: BLE001
                errors.append(e)

        if n_threads == 1:
            run(self._lanes[0])
        else:
            threads = [threading.Thread(target=run, args=(lane,)) for lane in self._lanes[:n_threads]]
            for t in threads:
                t.start()
            for t in threads:
                t.join()
        if errors:
            raise errors[0]

    def build_jk(self, TAs: Sequence[np.ndarray], MO_coeff_epsilon: float,
                 densities: Callable[[int], Sequence[np.ndarray]], indices: Sequence[int] | None = None,
                 ) -> dict[int, list[tuple[np.ndarray, np.ndarray]]]:
        """The all-fragment CONSUMER pipeline: every fragment is transformed with (ij|kl) left
        resident, ``scf.hf.dot_eri_dm``-type J/K contractions with the symmetric ``(nmo, nmo)`` matrices
        ``densities(i)`` are enqueued right behind it on the same stream, and only the ``nmo x nmo`` results travel
        to the host (pinned, asynchronous).  Nothing of size ``npair^2`` ever leaves HBM, so the host-memory bound of
        the full delivery does not apply.  Returns ``{i: [(vj, vk), ...]}``."""
        assert self.tensor is not None, "stage() first"
        idx = list(range(len(TAs))) if indices is None else list(indices)
        if not idx:
            return {}
        metric = None if self.tensor.whitened else self.metric
        # the densities of fragment f are formed on the host AFTER the kernels of f have been enqueued, i.e. while the
        # device works on f (and still has f-1's J/K queued): host preparation is hidden, not serialised in front
        first = [np.asfortranarray(d, dtype=np.float64) for d in densities(idx[0])]
        nd = len(first)
        total = sum(2 * nd * TAs[i].shape[1] ** 2 for i in idx)
        if self._jk_pool is None or self._jk_pool.size < total:
            if self._jk_pool is not None:
                self.ctx.synchronize()
                _lib.pinned_free(self._jk_pool)
            self._jk_pool = _lib.pinned_empty((max(total, 1),))
        pool = self._jk_pool
        views: dict[int, list[tuple[np.ndarray, np.ndarray]]] = {}
        off = 0
        try:
            for i in idx:
                nmo = TAs[i].shape[1]
                q.transform_integral_resident(self.tensor, TAs[i], self.screen, metric, MO_coeff_epsilon, ws=self.ws)
                self._note_borderline(i)
                dms = first if i == idx[0] else [np.asfortranarray(d, dtype=np.float64) for d in densities(i)]
                if len(dms) != nd:
                    raise ValueError("densities(i) must return the same number of matrices for every fragment")
                views[i] = []
                for d in dms:
                    vj = pool[off: off + nmo * nmo].reshape(nmo, nmo, order="F")
                    vk = pool[off + nmo * nmo: off + 2 * nmo * nmo].reshape(nmo, nmo, order="F")
                    off += 2 * nmo * nmo
                    self.ws.jk_enqueue(d, vj, vk)
                    views[i].append((vj, vk))
        finally:
            self.ctx.synchronize()
        return {i: [(np.array(vj), np.array(vk)) for vj, vk in views[i]] for i in idx}


class DistributedQueue:
    """Fragment indices handed out dynamically ACROSS RANKS (one process per GPU): an atomic counter in the process
    group's key-value store (``store.add``), so a rank whose PCIe link or host share is faster simply takes more
    fragments -- the fragment-parallel build has no other inter-rank traffic.  Every rank constructs it with the same
    ``order`` (largest first) and the same ``name``; iterate once."""

    def __init__(self, dist, order: Sequence[int], name: str, store=None):
        if store is None:
            from torch.distributed import distributed_c10d as c10d  # noqa: PLC0415

            store = c10d._get_default_store()
        self._store, self._order, self._key = store, list(order), f"qemb_fragment_queue/{name}"
        self._lock = threading.Lock()
        self.taken: list[int] = []

    def __iter__(self):
        return self

    def __next__(self) -> int:
        with self._lock:
            k = int(self._store.add(self._key, 1)) - 1
            if k >= len(self._order):
                raise StopIteration
            self.taken.append(self._order[k])
            return self._order[k]


class _SharedQueue:
    """Thread-safe iterator over fragment indices (largest first): devices and lanes pull from it dynamically."""

    def __init__(self, order):
        self._it, self._lock = iter(order), threading.Lock()

    def __iter__(self):
        return self

    def __next__(self):
        with self._lock:
            return next(self._it)


def build_all_fragments(unique_dense_data, shape, exch_reachable, L_PQ, S_abs, TAs, MO_coeff_epsilon, sink: Sink,
                        *, devices: Sequence[int] = (0,), whiten: bool = True,
                        tensors: Sequence[q.SemiSparseSym3DTensor] | None = None, n_threads: int = 1) -> None:
    """One process, one thread per device (x ``n_threads`` lanes each), fragments handed out dynamically (largest
    first) from one shared queue; every lane keeps its transform -> copy -> sink pipeline full.  ``tensors`` (one per
    device, already staged) replaces ``unique_dense_data``.  ``sink`` calls are serialised."""
    naux = shape[0]
    queue = _SharedQueue(sorted(range(len(TAs)), key=lambda i: -fragment_cost(naux, TAs[i].shape[1])))
    max_nmo = max(t.shape[1] for t in TAs)
    lock = threading.Lock()
    errors: list[BaseException] = []

    def locked_sink(i, G):
        with lock:
            return sink(i, G)

    def worker(k: int, dev: int):
        try:
            b = FragmentERIBuilder(dev)
            b.stage(unique_dense_data, shape, exch_reachable, L_PQ, S_abs, whiten=whiten,
                    tensor=None if tensors is None else tensors[k])
            b._build_from(queue, max_nmo, TAs, MO_coeff_epsilon, locked_sink, n_threads)
            b.close()
        except BaseException as e:  # noqa: BLE001
            errors.append(e)

    threads = [threading.Thread(target=worker, args=(k, d)) for k, d in enumerate(devices)]
    for t in threads:
        t.start()
    for t in threads:
        t.join()
    if errors:
        raise errors[0]


# ---------------------------------------------------------------------------
# QuEmb seam (needs quemb + pyscf in the environment)
# ---------------------------------------------------------------------------


def transform_sparse_DF_integral_b200(mf, Fobjs, auxbasis, file_eri_handler, AO_coeff_epsilon: float,
                                      MO_coeff_epsilon: float, n_threads: int = 1, precompute_P_mu_nu: bool = True,
                                      devices: Sequence[int] = (0,), async_write: bool = True) -> None:
    """Drop-in sibling of ``transform_sparse_DF_integral_gpu`` (``eri_sparse_DF.py:683-704``).

    Host-side one-offs that need libcint stay the reference's own functions (``approx_S_abs``, ``int2c2e``,
    ``_aux_e2`` = ``int3c2e`` blocks, SciPy Cholesky); the AO-AO screening rule and the block loop of
    ``get_sparse_P_mu_nu`` are restated in :mod:`quemb_b200.staging` and land the integrals directly in HBM.
    Writes dataset ``fobj.dname`` of shape ``(npair, npair)`` float64 per fragment (``eri_sparse_DF.py:630-631``);
    with ``async_write`` the ``create_dataset`` calls run on a background thread fed from the pinned output ring.
    ``n_threads`` keeps the reference's meaning (``ThreadPoolExecutor(n_threads)``, ``eri_sparse_DF.py:649-653``): that
    many host threads per device feed fragments concurrently, each on CUDA streams of its own."""
    from pyscf.ao2mo.addons import restore  # noqa: PLC0415
    from pyscf.df.addons import make_auxmol  # noqa: PLC0415
    from quemb.molbe import eri_sparse_DF as ref  # noqa: PLC0415
    from scipy.linalg import cholesky  # noqa: PLC0415

    from .eri_file import AsyncDatasetWriter  # noqa: PLC0415

    q.set_log_level(logging.getLogger().getEffectiveLevel())  # eri_sparse_DF.py:605

    mol = mf.mol
    auxmol = make_auxmol(mol, auxbasis=auxbasis)
    S_abs = ref.approx_S_abs(mol)
    PQ = auxmol.intor("int2c2e")
    L = cholesky(PQ, lower=True)
    nao, naux = int(mol.nao), int(auxmol.nao)
    TAs = [np.asfortranarray(f.TA) for f in Fobjs]

    writer = AsyncDatasetWriter(file_eri_handler) if async_write else None
    lock = threading.Lock()

    def sink(i: int, G: np.ndarray):
        eri = restore("4", G, Fobjs[i].TA.shape[1])  # identity for an (npair, npair) input
        if writer is not None:
            return writer.submit(Fobjs[i].dname, eri)
        with lock:  # h5py serialises writers anyway
            file_eri_handler.create_dataset(Fobjs[i].dname, data=eri)
        return None

    try:
        if precompute_P_mu_nu:
            exch = staging.get_AO_per_AO(S_abs, AO_coeff_epsilon, None)
            lists = staging.exch_lists(exch, nao)
            # libcint evaluates every block once; each device receives its own copy
            tensors = staging.get_sparse_P_mu_nu(mol, auxmol, exch, aux_e2=ref._aux_e2,
                                                 ctx=[q.default_context(d) for d in devices])
            build_all_fragments(None, (naux, nao, nao), lists, L, S_abs, TAs, MO_coeff_epsilon, sink, devices=devices,
                                whiten=True, tensors=tensors, n_threads=max(1, int(n_threads)))
        else:
            b = FragmentERIBuilder(devices[0])
            metric = q.GPU_MatrixHandle(L, ctx=b.ctx)  # uploaded and block-inverted once, not per fragment
            screen = q._ScreenHandle(np.asarray(S_abs), b.ctx)
            for i, fobj in enumerate(Fobjs):
                exch = staging.get_AO_per_AO(S_abs, AO_coeff_epsilon, fobj.TA)
                t = staging.get_sparse_P_mu_nu(mol, auxmol, exch, aux_e2=ref._aux_e2, ctx=b.ctx)
                b.stage(None, None, None, metric, screen, whiten=False, tensor=t)
                b.build(TAs, MO_coeff_epsilon, sink, indices=[i])
            b.close()
    finally:
        if writer is not None:
            writer.close()
