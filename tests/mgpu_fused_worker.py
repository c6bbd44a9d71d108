"""torchrun worker of test_fused_peer_accumulate_two_gpus: aux-sharded build with the reduce-sum fused into the K3
epilogue (system-scope red.add.f64 into the owner's matrix over NVLink peer memory), checked against the oracle on
the owner rank of every fragment.  Not collected by pytest."""
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
from quemb_b200.sharded import FusedAuxShardedBuilder

local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
rank, world = dist.get_rank(), dist.get_world_size()
p = syn.make_chain_problem(n_units=10, ao_per_unit=12, n_end=3, naux=333, frag_halfwidth=1, n_bath=20, sigma=0.7,
                           cond=1e8, seed=110)
b = FusedAuxShardedBuilder(dist, device=local)
b.stage(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable, p.L_PQ, p.S_abs)
errs, sym = {}, {}


def sink(i, G):
    ref = oracle.transform_integral(p.unique_dense_data, p.exch_reachable, p.TAs[i], p.S_abs, p.L_PQ,
                                    p.MO_coeff_epsilon)
    errs[i] = float(np.abs(G - ref).max())
    sym[i] = bool(np.array_equal(G, G.T))


owned = b.build(p.TAs, p.MO_coeff_epsilon, sink)
owned2 = b.build(p.TAs, p.MO_coeff_epsilon, sink)  # second pass reuses the exported matrices
out = [None] * world
dist.all_gather_object(out, {"rank": rank, "owned": owned, "errs": errs, "sym": sym, "same": owned == owned2})
if rank == 0:
    all_owned = sorted(i for o in out for i in o["owned"])
    worst = max(v for o in out for v in o["errs"].values())
    print(json.dumps(out))
    assert all_owned == list(range(len(p.TAs))), all_owned
    assert all(all(o["sym"].values()) for o in out) and all(o["same"] for o in out)
    assert worst <= 1e-10, worst
    print(f"FUSED_OK world {world} fragments {len(p.TAs)} max err {worst:.3e}")
dist.destroy_process_group()
