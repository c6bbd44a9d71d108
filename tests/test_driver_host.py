"""CPU: host-side logic of the multi-GPU path -- LPT partition of fragments over ranks, the
world_size-2 (gloo) plumbing the benchmark uses (barrier, max-over-ranks), and the reference arm's
rank-0-only behaviour."""
from __future__ import annotations

import json
import os
import socket
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

from quemb_b200.driver import fragment_cost, partition_lpt

ROOT = Path(__file__).resolve().parent.parent


def test_partition_lpt_covers_and_balances():
    rng = np.random.default_rng(0)
    nmos = list(rng.integers(100, 230, size=32))
    costs = [fragment_cost(2716, int(n)) for n in nmos]
    for world in (1, 2, 4, 8):
        parts = partition_lpt(costs, world)
        assert sorted(i for p in parts for i in p) == list(range(32))
        loads = [sum(costs[i] for i in p) for p in parts]
        # LPT guarantee: max load <= mean + largest item
        assert max(loads) <= sum(costs) / world + max(costs) + 1e-6
    assert partition_lpt([], 4) == [[], [], [], []]
    assert partition_lpt([1.0], 4) == [[0], [], [], []]


def test_chain_fragments_partition_is_deterministic():
    costs = [fragment_cost(2716, n) for n in [154, 201] + [225] * 28 + [201, 154]]
    a, b = partition_lpt(costs, 8), partition_lpt(costs, 8)
    assert a == b
    assert all(len(p) == 4 for p in a)


def _free_port() -> int:
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


WORKER = r"""
import os, sys, json
sys.path.insert(0, sys.argv[1])
import torch, torch.distributed as dist
from quemb_b200.driver import fragment_cost, partition_lpt
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
costs = [fragment_cost(700, n) for n in [110, 134, 158, 153, 153, 158, 134, 110]]
mine = partition_lpt(costs, world)[rank]
gathered = [None] * world
dist.all_gather_object(gathered, mine)
t = torch.tensor([float(sum(costs[i] for i in mine))], dtype=torch.float64)
dist.all_reduce(t, op=dist.ReduceOp.MAX)        # "max over ranks" of the per-rank time proxy
dist.barrier()
if rank == 0:
    print(json.dumps({"parts": gathered, "max_load": t.item(), "total": sum(costs)}))
dist.destroy_process_group()
"""


@pytest.mark.timeout(300)
def test_world_size_2_gloo_partition(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    port = _free_port()
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", LOCAL_RANK=str(r), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, str(script), str(ROOT)], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.PIPE, text=True))
    outs = [p.communicate(timeout=240) for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    res = json.loads(outs[0][0].strip().splitlines()[-1])
    assert sorted(i for part in res["parts"] for i in part) == list(range(8))
    assert set(res["parts"][0]).isdisjoint(res["parts"][1])
    assert res["total"] / 2 <= res["max_load"] <= 0.6 * res["total"]


@pytest.mark.timeout(600)
def test_reference_arm_prints_one_line_on_rank0_only():
    base = [sys.executable, str(ROOT / "bench.py"), "--impl", "reference", "--workload", "tiny", "--gpus", "2",
            "--steps", "1", "--warmup", "1"]
    r1 = subprocess.run(base, env=dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1"), capture_output=True,
                        text=True, timeout=300)
    assert r1.returncode == 0 and r1.stdout.strip() == ""
    r0 = subprocess.run(base, env=dict(os.environ, RANK="0", WORLD_SIZE="2", LOCAL_RANK="0"), capture_output=True,
                        text=True, timeout=300)
    assert r0.returncode == 0, r0.stderr
    line = json.loads(r0.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["metric"] == "all_fragment_eri_build_time"
    assert line["cpu_baseline"]["kind"] in ("reference", "port") and line["value"] > 0
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["higher_is_better"] is False


def test_aux_slabs_cover_and_align():
    from quemb_b200.sharded import aux_slabs, owner_of

    for naux in (16, 72, 700, 2716, 9024):
        for world in (1, 2, 3, 4, 8):
            slabs = aux_slabs(naux, world)
            assert slabs[0][0] == 0 and slabs[-1][1] == naux and len(slabs) == world
            for (a0, a1), (b0, b1) in zip(slabs, slabs[1:]):
                assert a1 == b0 and a0 <= a1
            assert all(a % 16 == 0 for a, _ in slabs)
            sizes = [b - a for a, b in slabs]
            if naux >= 16 * world:
                assert min(sizes) > 0 and max(sizes) - min(sizes) <= 16 + naux % 16
    assert [owner_of(k, 4) for k in range(6)] == [0, 1, 2, 3, 0, 1]
    with pytest.raises(ValueError):
        aux_slabs(0, 2)


SHARDED_WORKER = r"""
import os, sys, json
sys.path.insert(0, sys.argv[1])
import numpy as np, scipy.linalg as sla
import torch, torch.distributed as dist
from oracle import oracle                      # CPU checker standing in for the device hooks
from quemb_b200 import synthetic as syn
from quemb_b200.sharded import AuxShardedBuilder, aux_slabs

class CpuSlabBuilder(AuxShardedBuilder):
    # the orchestration under test is the base class; only the three device hooks are replaced
    def _stage_slab(self, M, shape, exch, L, S_abs):
        p0, p1 = self.slab
        W = sla.solve_triangular(L, M, lower=True)
        self.W = np.asfortranarray(W[p0:p1]); self.exch = exch; self.S_abs = S_abs
    def _partial(self, k, TA, eps):
        G = oracle.transform_integral(self.W, self.exch, TA, self.S_abs, np.eye(self.W.shape[0]), eps)
        return torch.from_numpy(np.ascontiguousarray(G).reshape(-1).copy())
    def _to_host(self, G, npair):
        return G.numpy().reshape(npair, npair)
    def _collective_done(self):
        pass
    def _before_collective(self):
        pass

dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
p = syn.make_chain_problem(n_units=6, ao_per_unit=6, n_end=1, naux=88, frag_halfwidth=1, n_bath=6, sigma=0.5, seed=5)
b = AuxShardedBuilder.__new__(CpuSlabBuilder); AuxShardedBuilder.__init__(b, dist, device=0)
b.stage(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable, p.L_PQ, p.S_abs)
got = {}
owned = b.build(p.TAs, p.MO_coeff_epsilon, lambda i, G: got.__setitem__(i, G.copy()))
errs = {}
for i, G in got.items():
    ref = oracle.transform_integral(p.unique_dense_data, p.exch_reachable, p.TAs[i], p.S_abs, p.L_PQ, p.MO_coeff_epsilon)
    errs[i] = float(np.abs(G - ref).max())
out = [None] * world
dist.all_gather_object(out, {"rank": rank, "slab": b.slab, "owned": owned, "errs": errs})
if rank == 0:
    print(json.dumps({"ranks": out, "nfrag": len(p.TAs), "naux": p.naux}))
dist.destroy_process_group()
"""


@pytest.mark.timeout(300)
def test_world_size_2_gloo_aux_sharded_reduce(tmp_path):
    """N>1 path of the aux-sharded build on CPU: slabs, rotating owners, async reduce-sum, sink on the owner."""
    script = tmp_path / "sharded_worker.py"
    script.write_text(SHARDED_WORKER)
    port = _free_port()
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", LOCAL_RANK=str(r), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port), OMP_NUM_THREADS="2")
        procs.append(subprocess.Popen([sys.executable, str(script), str(ROOT)], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.PIPE, text=True))
    outs = [p.communicate(timeout=240) for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    res = json.loads(outs[0][0].strip().splitlines()[-1])
    r0, r1 = res["ranks"]
    assert r0["slab"][0] == 0 and r0["slab"][1] == r1["slab"][0] and r1["slab"][1] == res["naux"]
    assert sorted(r0["owned"] + r1["owned"]) == list(range(res["nfrag"]))
    assert r0["owned"] == list(range(0, res["nfrag"], 2)) and r1["owned"] == list(range(1, res["nfrag"], 2))
    errs = {**r0["errs"], **r1["errs"]}
    assert len(errs) == res["nfrag"] and max(errs.values()) <= 1e-10, errs


def test_eri_file_container_roundtrip(tmp_path):
    """Same dataset names / shape / dtype contract as the reference's HDF5 file (eri_sparse_DF.py:631)."""
    from quemb_b200.eri_file import NpyDirFile, open_eri_file

    rng = np.random.default_rng(0)
    G = rng.standard_normal((21, 21))
    G = G + G.T
    with open_eri_file(tmp_path / "eri_file.h5", "w") as f:
        f.create_dataset("f0", data=G)
        f.create_dataset("f1", data=np.asfortranarray(2 * G))
    with open_eri_file(tmp_path / "eri_file.h5", "r") as f:
        assert np.array_equal(np.array(f.get("f0")), G) and np.array_equal(np.array(f["f1"]), 2 * G)
        assert np.array(f.get("f0")).dtype == np.float64 and np.array(f.get("f0")).shape == (21, 21)
        if isinstance(f, NpyDirFile):
            assert f.get("f7") is None and f.keys() == ["f0", "f1"]
            with pytest.raises(OSError):
                f.create_dataset("f2", data=G)


FUSED_WORKER = r"""
import os, sys, json, fcntl
sys.path.insert(0, sys.argv[1])
import numpy as np, scipy.linalg as sla
import torch.distributed as dist
from oracle import oracle                      # CPU checker standing in for the device hooks
from quemb_b200 import synthetic as syn
from quemb_b200.sharded import FusedAuxShardedBuilder

SCRATCH = sys.argv[2]

class CpuFused(FusedAuxShardedBuilder):
    # peer memory = a file both processes map; the K3 atomics = read-modify-write under a file lock
    def _stage_slab(self, M, shape, exch, L, S_abs):
        p0, p1 = self.slab
        self.W = np.asfortranarray(sla.solve_triangular(L, M, lower=True)[p0:p1]); self.exch = exch; self.S_abs = S_abs
    def _alloc(self, nbytes):
        path = os.path.join(SCRATCH, f"g_{self.rank}_{len(os.listdir(SCRATCH))}.bin")
        m = np.memmap(path, dtype=np.float64, mode="w+", shape=(nbytes // 8,))
        return m, path.encode()
    def _open(self, handle):
        return np.memmap(handle.decode(), dtype=np.float64, mode="r+")
    def _zero(self, target, npair):
        target[: npair * npair] = 0.0; target.flush()
    def _push_partial(self, TA, eps, target, npair):
        G = oracle.transform_integral(self.W, self.exch, TA, self.S_abs, np.eye(self.W.shape[0]), eps)
        low = np.tril(np.ones((npair, npair), dtype=bool))            # lower tiles only, like the kernel
        with open(os.path.join(SCRATCH, "lock"), "w") as lk:
            fcntl.flock(lk, fcntl.LOCK_EX)
            v = np.asarray(target[: npair * npair]).reshape(npair, npair).copy()
            v[low] += np.asarray(G)[low]
            target[: npair * npair] = v.reshape(-1); target.flush()
            fcntl.flock(lk, fcntl.LOCK_UN)
    def _owner_finish(self, target, npair, to_host):
        G = np.array(target[: npair * npair]).reshape(npair, npair)
        G = np.tril(G) + np.tril(G, -1).T
        return G

dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
p = syn.make_chain_problem(n_units=6, ao_per_unit=6, n_end=1, naux=88, frag_halfwidth=1, n_bath=6, sigma=0.5, seed=6)
b = CpuFused(dist, device=0)
b.stage(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable, p.L_PQ, p.S_abs)
errs = {}
def sink(i, G):
    ref = oracle.transform_integral(p.unique_dense_data, p.exch_reachable, p.TAs[i], p.S_abs, p.L_PQ, p.MO_coeff_epsilon)
    errs[i] = float(np.abs(G - ref).max())
owned = b.build(p.TAs, p.MO_coeff_epsilon, sink)
out = [None] * world
dist.all_gather_object(out, {"rank": rank, "owned": owned, "errs": errs})
if rank == 0:
    print(json.dumps({"ranks": out, "nfrag": len(p.TAs)}))
dist.destroy_process_group()
"""


@pytest.mark.timeout(300)
def test_world_size_2_gloo_fused_peer_accumulate(tmp_path):
    """Orchestration of the fused K3 -> peer-memory reduction (zero / barrier / accumulate / barrier / mirror, two
    exported matrices per rank, rotating owners) with file-backed shared memory standing in for NVLink."""
    script = tmp_path / "fused_worker.py"
    script.write_text(FUSED_WORKER)
    scratch = tmp_path / "shm"
    scratch.mkdir()
    port = _free_port()
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", LOCAL_RANK=str(r), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port), OMP_NUM_THREADS="2")
        procs.append(subprocess.Popen([sys.executable, str(script), str(ROOT), str(scratch)], env=env,
                                      stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    outs = [p.communicate(timeout=240) for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    res = json.loads(outs[0][0].strip().splitlines()[-1])
    r0, r1 = res["ranks"]
    assert sorted(r0["owned"] + r1["owned"]) == list(range(res["nfrag"]))
    errs = {**r0["errs"], **r1["errs"]}
    assert len(errs) == res["nfrag"] and max(errs.values()) <= 1e-10, errs


# ---------------------------------------------------------------------------------------------------------------------
# the fragment loop itself (lanes, shared queue, buffer hand-back) with the device replaced by a host fake
# ---------------------------------------------------------------------------------------------------------------------
class _FakeDevice:
    """Stands in for quemb_b200.eri_sparse_DF inside quemb_b200.driver: transforms write a recognisable matrix into
    the caller's pinned view, tickets complete in order.  Only the HOST logic of driver.py is under test here."""

    def __init__(self):
        import threading

        self.lock = threading.Lock()
        self.calls = []

    class Ctx:
        device = 0

        def synchronize(self):
            pass

    class WS:
        def __init__(self, ctx=None):
            self.last_ticket = 0
            self.borderline = 0
            self.waited = []

        def wait_output(self, ticket):
            self.waited.append(ticket)

    class Tensor:
        whitened = True

    def install(self, monkeypatch):
        from quemb_b200 import _lib, driver

        fake = self
        monkeypatch.setattr(driver.q, "default_context", lambda d=None: _FakeDevice.Ctx())
        monkeypatch.setattr(driver.q, "Workspace", _FakeDevice.WS)
        monkeypatch.setattr(driver.q, "output_ring_depth", lambda: 3)
        monkeypatch.setattr(_lib, "pinned_empty", lambda shape, dtype=np.float64: np.empty(shape, dtype=dtype))
        monkeypatch.setattr(_lib, "pinned_free", lambda a: None)

        def transform_integral(tensor, TA, screen, metric, eps, ws=None, out=None, synchronize=True):
            out[...] = float(TA[0, 0])
            ws.last_ticket += 1
            with fake.lock:
                fake.calls.append((id(ws), int(TA[0, 0])))
            return out

        monkeypatch.setattr(driver.q, "transform_integral", transform_integral)


@pytest.mark.parametrize("n_threads", [1, 3])
def test_fragment_loop_delivers_every_fragment_once(monkeypatch, n_threads):
    from concurrent.futures import ThreadPoolExecutor

    from quemb_b200 import driver

    fake = _FakeDevice()
    fake.install(monkeypatch)
    TAs = [np.full((4, 2 + (i % 3)), float(i)) for i in range(11)]  # fragment i is recognisable by TA[0, 0] = i
    b = driver.FragmentERIBuilder(0, n_host_buffers=2)
    b.tensor, b.metric, b.screen = _FakeDevice.Tensor(), None, None
    seen = {}
    pool = ThreadPoolExecutor(max_workers=1)

    def sink(i, G):  # returns a Future: the buffer must stay untouched until it is done
        def check():
            import time

            time.sleep(0.002)
            npair = TAs[i].shape[1] * (TAs[i].shape[1] + 1) // 2
            assert G.shape == (npair, npair) and np.all(G == float(i)), (i, G)
            seen[i] = seen.get(i, 0) + 1
            return i

        return pool.submit(check)

    b.build(TAs, 1e-5, sink, n_threads=n_threads)
    pool.shutdown()
    assert seen == {i: 1 for i in range(len(TAs))}
    assert len({w for w, _ in fake.calls}) == n_threads  # one workspace per lane
    assert len(b._lanes) >= n_threads and all(len(l.bufs) == 2 for l in b._lanes[:n_threads])
    b.close()


def test_build_all_fragments_shared_queue_largest_first(monkeypatch):
    from quemb_b200 import driver

    fake = _FakeDevice()
    fake.install(monkeypatch)
    monkeypatch.setattr(driver.FragmentERIBuilder, "stage", lambda self, *a, **k: (setattr(self, "tensor", _FakeDevice.Tensor())))
    TAs = [np.full((4, n), float(i)) for i, n in enumerate([2, 5, 3, 6, 4, 2, 7])]
    got = []
    driver.build_all_fragments(None, (8, 4, 4), None, None, None, TAs, 1e-5, lambda i, G: got.append(i),
                               devices=(0, 1), n_threads=2)
    assert sorted(got) == list(range(len(TAs)))
    order = [i for _, i in fake.calls]
    assert order[0] == 6 and set(order[:2]) == {6, 3}  # largest fragments are handed out first
    with pytest.raises(ValueError):
        driver.FragmentERIBuilder(0, n_host_buffers=0)


QUEUE_WORKER = r"""
import os, sys, json, time
sys.path.insert(0, sys.argv[1])
import torch.distributed as dist
from quemb_b200.driver import DistributedQueue
dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
order = [7, 3, 5, 0, 1, 6, 2, 4, 8]
got = {}
for name in ("a", "b"):
    q = DistributedQueue(dist, order, name)
    mine = []
    for i in q:
        mine.append(i)
        time.sleep(0.002 * (1 + 3 * rank))      # rank 1 is slower: it must end up with fewer fragments
    got[name] = mine
    assert q.taken == mine
    dist.barrier()
out = [None] * world
dist.all_gather_object(out, got)
if rank == 0:
    print(json.dumps(out))
dist.destroy_process_group()
"""


@pytest.mark.timeout(300)
def test_world_size_2_gloo_distributed_fragment_queue(tmp_path):
    """Cross-rank dynamic hand-out (atomic counter in the process group's store): every fragment exactly once, in the
    given order, the faster rank takes more; a second queue under another name starts from scratch."""
    script = tmp_path / "queue_worker.py"
    script.write_text(QUEUE_WORKER)
    port = _free_port()
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", LOCAL_RANK=str(r), MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, str(script), str(ROOT)], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.PIPE, text=True))
    outs = [p.communicate(timeout=240) for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    r0, r1 = json.loads(outs[0][0].strip().splitlines()[-1])
    for name in ("a", "b"):
        assert sorted(r0[name] + r1[name]) == list(range(9))
        assert len(r0[name]) > len(r1[name])
        merged = [7, 3, 5, 0, 1, 6, 2, 4, 8]
        assert [i for i in merged if i in r0[name]] == r0[name]  # each rank sees its share in queue order


# ---- output-mode autotune: the tie-break rule (host only) ---------------------------------------------------------------
# (box, P0, P, M, trial seconds [full squares, triangle, split], split fraction, mode that won the BUILD on that box)
_AUTOTUNE_RECORDS = [
    # profiles/bench_r2l_n1_after_noalloc.json: triangle 12 % ahead in the trials; build 2.09 s
    ("n1 fast-pcie", 57.22, 51.96, 54.08, [0.037860, 0.033312, -1.0], 0.0, 1),
    # profiles/bench_r2i_n1_autotune_trials.json
    ("n1 fast-pcie b", 57.22, 48.32, 62.65, [0.037924, 0.034501, -1.0], 0.0, 1),
    # profiles/bench_r2k_n1_group256.json / bench_r2j_n1_final.json: PCIe drops to 32-34 GB/s under the mirror; triangle
    # builds 2.30 s, full squares 2.81 s
    ("n1 slow-pcie a", 55.66, 32.23, 65.67, [0.039073, 0.037569, -1.0], 0.0, 1),
    ("n1 slow-pcie b", 55.53, 33.99, 65.68, [0.040409, 0.039709, -1.0], 0.0, 1),
    # profiles/bench_r2m_n1_trial_tie.json: the trials TIE with full squares 0.5 % ahead -> 2.91 s instead of 2.30 s
    ("n1 slow-pcie tie", 54.92, 32.73, 63.53, [0.038093, 0.038290, -1.0], 0.0, 1),
    # profiles/bench_r2i_n2_autotune_trials.json, both ranks: the model alone wanted the split, which is 11 % behind in
    # the trials; full squares build 1.41 s against 1.97 s for the triangle
    ("n2 rank0", 56.95, 41.52, 25.62, [0.038626, 0.039250, 0.043005], 0.237, 0),
    ("n2 rank1", 56.77, 42.22, 23.42, [0.038131, 0.040757, 0.043655], 0.287, 0),
]


@pytest.mark.parametrize("rec", _AUTOTUNE_RECORDS, ids=[r[0] for r in _AUTOTUNE_RECORDS])
def test_autotune_pick_chooses_the_mode_that_won_the_build(rec):
    import ctypes

    from quemb_b200 import _lib

    _, p0, p, m, secs, f, want = rec
    pick = ctypes.c_int(-7)
    a, u = (ctypes.c_double * 3)(*secs), (ctypes.c_double * 3)(1.0, 0.0, f)
    _lib.check(_lib.lib().qemb_autotune_pick(a, u, p0, p, m, ctypes.byref(pick)))
    assert pick.value == want


def test_autotune_pick_edge_cases():
    import ctypes

    from quemb_b200 import _lib

    def pick(secs, f=0.0, rates=(55.0, 40.0, 40.0)):
        out = ctypes.c_int(-7)
        _lib.check(_lib.lib().qemb_autotune_pick((ctypes.c_double * 3)(*secs), (ctypes.c_double * 3)(1.0, 0.0, f), *rates,
                                                 ctypes.byref(out)))
        return out.value

    assert pick([-1.0, -1.0, -1.0]) == -1                      # nothing ran: the caller keeps the model's choice
    assert pick([0.05, -1.0, -1.0]) == 0                       # a single candidate
    assert pick([0.030, 0.040, -1.0]) == 0                     # decisive trials beat the model (which prefers 1 here)
    assert pick([0.040, 0.030, 0.0301], f=0.3, rates=(55.0, 40.0, 20.0)) == 2  # tie 1 / 2: the model prefers the split
    assert pick([0.040, 0.041, -1.0], rates=(0.0, 0.0, 0.0)) == 0  # no rates: fastest trial
