"""Aux-sharded fragments: the auxiliary index P of ONE fragment is split over the ranks of a process group.

BASELINE.json north_star: "only fragments whose B tensor exceeds one GPU shard the auxiliary index P and
combine with an NCCL reduce-sum over NVLink" ("Oversize fragments").  The reference has no
counterpart: its worker holds all of (P|mu nu) and (P|ij) in one address space
(``src/quemb/molbe/eri_sparse_DF.py:621-631``).

K1, K2a and K3 never mix auxiliary rows once the metric has been applied to (P|mu nu) (design option O1),

    (ij|kl) = sum_P X[P,ij] X[P,kl] = sum_r  sum_{P in slab r} X[P,ij] X[P,kl],     X = (L^-1 M) o TA o TA,

so rank r keeps rows ``[p_r, p_{r+1})`` of ``L^-1 (P|mu nu)`` (1/world of the tensor, staged through
``SemiSparseSym3DTensor.whitened_slab``), runs the unchanged fragment kernels on its slab and the partial
Gram matrices are summed onto the fragment's OWNER rank with one ``reduce`` per fragment (NCCL over
NVLink/NVSwitch on the GPU box; owners rotate over the ranks so the device->host copy and the sink work of
consecutive fragments land on different ranks).  The screening lists A(i) are computed redundantly by every
rank from the same S_abs and TA with the same deterministic kernel, so every rank plans the same columns.

The sum order differs from the single-GPU run: results agree to the 1e-10 budget, not bit for bit.
"""

from __future__ import annotations

from typing import Callable, Sequence

import numpy as np

Sink = Callable[[int, np.ndarray], None]


def aux_slabs(naux: int, world: int, align: int = 16) -> list[tuple[int, int]]:
    """Split ``[0, naux)`` into ``world`` contiguous slabs whose interior bounds are multiples of ``align``
    (the kernels' 16-row k tile).  Slabs differ by at most one tile; trailing slabs may be empty when
    ``naux < align * world``."""
    if naux <= 0 or world <= 0:
        raise ValueError("naux and world must be positive")
    ntile = -(-naux // align)
    bounds = [min(naux, align * ((ntile * r) // world)) for r in range(world)] + [naux]
    return [(bounds[r], bounds[r + 1]) for r in range(world)]


def owner_of(k: int, world: int) -> int:
    """Rank that receives the summed (ij|kl) of the k-th fragment of a sharded build."""
    return k % world


class AuxShardedBuilder:
    """One rank's part of an aux-sharded all-fragment build.

    ``dist`` is ``torch.distributed`` (initialised by the caller, backend ``nccl`` on GPUs); ``group`` the
    process group the auxiliary index is split over.  The three hooks ``_stage_slab`` / ``_partial`` /
    ``_to_host`` are the device side; the CPU tests replace them to exercise the orchestration over gloo.
    """

    def __init__(self, dist, device: int = 0, group=None):
        self.dist = dist
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        self.device = device
        self.slab: tuple[int, int] | None = None
        self.tensor = None
        self.screen = None
        self.ws = None
        self.ctx = None
        self._G: list = [None, None]

    # -- device side ------------------------------------------------------------------------------------
    def _stage_slab(self, unique_dense_data, shape, exch_reachable, L_PQ, S_abs) -> None:
        from . import eri_sparse_DF as q

        self.ctx = q.default_context(self.device)
        self.ws = q.Workspace(self.ctx)
        metric = q.GPU_MatrixHandle(L_PQ, ctx=self.ctx)
        self.tensor = q.SemiSparseSym3DTensor.whitened_slab(unique_dense_data, shape, exch_reachable, metric,
                                                            self.slab, ctx=self.ctx)
        del metric  # L is not needed after staging: the slab is already whitened
        self.screen = q._ScreenHandle(np.asarray(S_abs), self.ctx)

    def _buffer(self, k: int, n: int):
        import torch

        b = self._G[k & 1]
        if b is None or b.numel() < n:
            b = torch.empty(n, dtype=torch.float64, device=torch.device("cuda", self.device))
            self._G[k & 1] = b
        return b[:n]

    def _partial(self, k: int, TA: np.ndarray, eps: float):
        """Partial (ij|kl) of this rank's slab, as a flat device tensor of npair^2 doubles."""
        from . import eri_sparse_DF as q

        nmo = TA.shape[1]
        npair = nmo * (nmo + 1) // 2
        G = self._buffer(k, npair * npair)
        q.transform_integral_device(self.tensor, TA, self.screen, None, eps, G.data_ptr(), npair, ws=self.ws)
        return G  # enqueued only: _before_collective orders torch's stream behind it, no host synchronisation

    def _torch_stream(self) -> int:
        import torch

        return torch.cuda.current_stream(self.device).cuda_stream

    def _before_collective(self) -> None:
        """The collective is enqueued on torch's stream: make that stream wait (on the device) for the kernels the
        library has enqueued so far -- a CUDA event, not the per-fragment ``ctx.synchronize()`` of round 1."""
        self.ctx.stream_handoff(self._torch_stream(), 0)

    def _to_host(self, G, npair: int) -> np.ndarray:
        return G.cpu().numpy().reshape(npair, npair)

    def _collective_done(self) -> None:
        """NCCL's ``work.wait()`` orders torch's stream after the collective; the library writes the next partial
        into the same buffer on its own stream, so that stream is made to wait for torch's (event, no host sync)."""
        self.ctx.stream_handoff(self._torch_stream(), 1)

    # -- orchestration ----------------------------------------------------------------------------------
    def stage(self, unique_dense_data, shape, exch_reachable, L_PQ, S_abs) -> None:
        naux = int(shape[0])
        self.slab = aux_slabs(naux, self.world)[self.rank]
        if self.slab[1] <= self.slab[0]:
            raise ValueError(f"rank {self.rank}: empty auxiliary slab (naux {naux} over {self.world} ranks)")
        self._stage_slab(unique_dense_data, shape, exch_reachable, L_PQ, S_abs)

    def build(self, TAs: Sequence[np.ndarray], MO_coeff_epsilon: float, sink: Sink | None = None,
              indices: Sequence[int] | None = None) -> list[int]:
        """Every rank calls this with the same arguments.  ``sink(i, G)`` runs on the owner rank of
        fragment i only; returns the fragments this rank owned."""
        idx = list(range(len(TAs))) if indices is None else list(indices)
        owned: list[int] = []
        pending = None  # (work handle, k, i, G, npair) of the previous fragment: its reduce overlaps our kernels
        for k, i in enumerate(idx):
            TA = np.asfortranarray(TAs[i], dtype=np.float64)
            nmo = TA.shape[1]
            npair = nmo * (nmo + 1) // 2
            G = self._partial(k, TA, MO_coeff_epsilon)
            if pending is not None:
                self._finish(pending, sink, owned)
            self._before_collective()
            work = self.dist.reduce(G, dst=self._global_rank(owner_of(k, self.world)), op=self.dist.ReduceOp.SUM,
                                    group=self.group, async_op=True)
            pending = (work, k, i, G, npair)
        if pending is not None:
            self._finish(pending, sink, owned)
        return owned

    def _global_rank(self, group_rank: int) -> int:
        if self.group is None:
            return group_rank
        return self.dist.get_global_rank(self.group, group_rank)

    def _finish(self, pending, sink: Sink | None, owned: list[int]) -> None:
        work, k, i, G, npair = pending
        work.wait()
        if owner_of(k, self.world) == self.rank:
            owned.append(i)
            if sink is not None:
                sink(i, self._to_host(G, npair))  # reads on torch's stream, behind the collective
        self._collective_done()


class FusedAuxShardedBuilder(AuxShardedBuilder):
    """Same aux-sharded build, but the reduce-sum is FUSED into K3: every rank's K3 epilogue adds its lower tiles
    straight into the owner rank's matrix through NVLink peer memory (``qemb_fragment_transform`` with
    ``out_is_device = 2``: system-scope ``red.add.f64`` from the tile epilogue), so the transfer of a tile overlaps
    the DMMAs of the following tiles and no separate collective (and no second npair^2 buffer per rank) is needed.

    Per fragment: the owner zero-fills one of its two exported matrices, a barrier publishes that, every rank runs
    screen / K1 / K2a / K3-accumulate on its slab with the owner's matrix as the target, a second barrier tells
    the owner that all contributions have landed, and the owner rebuilds the upper triangle on the device
    (``qemb_device_mirror_lower``).  The matrices are plain ``cudaMalloc`` buffers shared with ``cudaIpc*`` handles,
    exchanged once through ``all_gather_object``.  The sum order inside a tile entry depends on arrival order for
    more than two ranks (atomics): results agree with the unsharded path to the 1e-10 budget, not bit for bit, and
    are NOT run-to-run reproducible beyond two ranks.  This variant is therefore OPT-IN; the default and the one
    ``bench.py`` reports is :class:`AuxShardedBuilder` (NCCL ``reduce``, fixed reduction tree, deterministic).
    """

    def __init__(self, dist, device: int = 0, group=None):
        super().__init__(dist, device, group)
        self._own: list = [None, None]          # this rank's two matrices
        self._peer: list[list] = []             # [rank][0|1] -> target usable from this process
        self._cap = 0

    # -- device side (replaced by the CPU tests) -----------------------------------------------------------
    def _alloc(self, nbytes: int):
        import ctypes

        from ._lib import c_i64, c_p, check, lib

        dev = c_p()
        handle = (ctypes.c_ubyte * 64)()
        check(lib().qemb_ipc_alloc(self.ctx._h, c_i64(nbytes), ctypes.byref(dev), handle))
        return dev, bytes(handle)

    def _open(self, handle: bytes):
        import ctypes

        from ._lib import c_p, check, lib

        dev = c_p()
        buf = (ctypes.c_ubyte * 64).from_buffer_copy(handle)
        check(lib().qemb_ipc_open(self.ctx._h, buf, ctypes.byref(dev)))
        return dev

    def _zero(self, target, npair: int) -> None:
        from ._lib import c_i64, check, lib

        check(lib().qemb_device_zero(self.ctx._h, target, c_i64(8 * npair * npair)))
        self.ctx.synchronize()

    def _push_partial(self, TA: np.ndarray, eps: float, target, npair: int) -> None:
        """screen -> K1 -> K2a -> K3 on this rank's slab; K3 ADDS its lower tiles into ``target`` (maybe remote)."""
        import ctypes

        from . import eri_sparse_DF as q
        from ._lib import c_dbl, c_i64, c_int, check, lib, ptr

        check(lib().qemb_fragment_transform(
            self.ws._h, self.tensor._h, None, self.screen._h, ptr(TA), c_i64(q._ld(TA)), c_i64(TA.shape[0]),
            c_i64(TA.shape[1]), c_dbl(eps), c_i64(0), c_i64(-1), ctypes.cast(target, ctypes.c_void_p), c_i64(npair),
            c_int(2), None))
        self.ctx.synchronize()

    def _owner_finish(self, target, npair: int, to_host: bool):
        from ._lib import c_i64, check, lib, ptr

        check(lib().qemb_device_mirror_lower(self.ctx._h, target, c_i64(npair), c_i64(npair)))
        if not to_host:
            self.ctx.synchronize()
            return None
        out = np.empty((npair, npair), order="F")
        check(lib().qemb_device_to_host(self.ctx._h, ptr(out), target, c_i64(8 * npair * npair)))
        return out

    def _barrier(self) -> None:
        self.dist.barrier(group=self.group)

    # -- orchestration -----------------------------------------------------------------------------------------
    def _ensure_targets(self, max_npair: int) -> None:
        need = 8 * max_npair * max_npair
        if self._cap >= need:
            return
        if self._cap:
            raise RuntimeError("FusedAuxShardedBuilder: call build() with the largest fragments first or reserve() "
                               "once; exported matrices are not re-allocated")
        mine = [self._alloc(need) for _ in range(2)]
        self._own = [m[0] for m in mine]
        handles = [None] * self.world
        self.dist.all_gather_object(handles, [m[1] for m in mine], group=self.group)
        self._peer = []
        for rk in range(self.world):
            self._peer.append(list(self._own) if rk == self.rank else [self._open(h) for h in handles[rk]])
        self._cap = need

    def reserve(self, max_nmo: int) -> None:
        """Collective: allocate and exchange the exported matrices for fragments up to ``max_nmo`` orbitals."""
        self._ensure_targets(max_nmo * (max_nmo + 1) // 2)

    def build(self, TAs: Sequence[np.ndarray], MO_coeff_epsilon: float, sink: Sink | None = None,
              indices: Sequence[int] | None = None) -> list[int]:
        idx = list(range(len(TAs))) if indices is None else list(indices)
        if not idx:
            return []
        self.reserve(max(TAs[i].shape[1] for i in idx))
        owned: list[int] = []
        for k, i in enumerate(idx):
            TA = np.asfortranarray(TAs[i], dtype=np.float64)
            nmo = TA.shape[1]
            npair = nmo * (nmo + 1) // 2
            owner = owner_of(k, self.world)
            target = self._peer[owner][(k // self.world) & 1]
            if owner == self.rank:
                self._zero(target, npair)
            self._barrier()  # the owner's matrix is zero and nobody still reads its previous content
            self._push_partial(TA, MO_coeff_epsilon, target, npair)
            self._barrier()  # every rank's tiles have been added
            if owner == self.rank:
                owned.append(i)
                G = self._owner_finish(target, npair, sink is not None)
                if sink is not None:
                    sink(i, G)
        return owned
