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
