"""torchrun worker of test_aux_sharded_build_two_gpus: aux-sharded build over all ranks (NCCL reduce-sum),
checked against the oracle on the owner rank of every fragment.  Not collected by pytest."""
import json
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch
import torch.distributed as dist

from oracle import oracle
from quemb_b200 import synthetic as syn
from quemb_b200.sharded import AuxShardedBuilder

local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
rank, world = dist.get_rank(), dist.get_world_size()
p = syn.make_chain_problem(n_units=10, ao_per_unit=12, n_end=3, naux=333, frag_halfwidth=1, n_bath=20, sigma=0.7,
                           cond=1e8, seed=108)
b = AuxShardedBuilder(dist, device=local)
b.stage(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable, p.L_PQ, p.S_abs)
assert b.tensor.shape[0] == b.slab[1] - b.slab[0] < p.naux
errs = {}


def sink(i, G):
    ref = oracle.transform_integral(p.unique_dense_data, p.exch_reachable, p.TAs[i], p.S_abs, p.L_PQ,
                                    p.MO_coeff_epsilon)
    errs[i] = float(np.abs(G - ref).max())


owned = b.build(p.TAs, p.MO_coeff_epsilon, sink)

# column-split metric application with NCCL broadcasts == the single-rank one, bit for bit (same kernel per column)
from quemb_b200 import eri_sparse_DF as q
Ta = q.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable, ctx=b.ctx)
Tb = q.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable, ctx=b.ctx)
Ta.whiten(p.L_PQ)
Tb.whiten_distributed(p.L_PQ, dist)
assert Tb.whitened and np.array_equal(Ta.unique_dense_data, Tb.unique_dense_data), "whiten_distributed differs"
out = [None] * world
dist.all_gather_object(out, {"rank": rank, "owned": owned, "errs": errs, "slab": b.slab})
if rank == 0:
    all_owned = sorted(i for o in out for i in o["owned"])
    worst = max(v for o in out for v in o["errs"].values())
    print(json.dumps(out))
    assert all_owned == list(range(len(p.TAs))), all_owned
    assert worst <= 1e-10, worst
    print(f"SHARDED_OK world {world} fragments {len(p.TAs)} max err {worst:.3e}")
dist.destroy_process_group()
