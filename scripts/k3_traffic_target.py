"""ncu target for roofline.traffic: ONE all-fragment device build of the bench workload (C32H66-like, 32 fragments),
so that `ncu -k regex:gram_kernel --metrics dram__bytes_read.sum,dram__bytes_write.sum` sees exactly the 32 K3 launches
bench.py averages over.  scripts/k3_traffic_summarise.py turns the CSV into profiles/k3_traffic_r2.json."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch

import bench
from quemb_b200 import eri_sparse_DF as q
from quemb_b200.driver import FragmentERIBuilder

p = bench.make_workload("c32h66", use_gpu_matmul=True)
b = FragmentERIBuilder(0)
T = q.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable, ctx=b.ctx)
b.stage(None, None, None, p.L_PQ, p.S_abs, whiten=True, tensor=T)
n = max(t.shape[1] for t in p.TAs)
npair = n * (n + 1) // 2
G = torch.empty(npair * npair, dtype=torch.float64, device="cuda")
b.build_device(p.TAs, p.MO_coeff_epsilon, G.data_ptr())
b.ctx.synchronize()
print("built", len(p.TAs), "fragments")
