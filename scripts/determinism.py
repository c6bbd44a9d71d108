"""Development aid: run-to-run bitwise reproducibility of K1 / K2a / K3 at the C32H66-like size."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import bench
from quemb_b200 import eri_sparse_DF as q

p = bench.make_workload("c32h66", use_gpu_matmul=True)
ctx = q.default_context(0)
T = q.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable, ctx=ctx)
for f in (1, 2, 17):
    TA = p.TAs[f]
    A = q.get_AO_per_MO(TA, p.S_abs, p.MO_coeff_epsilon)
    Bs, Ss = [], []
    for rep in range(3):
        B = q.contract_with_TA_1st(TA, T, A)
        Bd = B.dense_data
        sym = q.contract_with_TA_2nd_to_sym_dense(B, TA)
        Bs.append(Bd); Ss.append(sym)
    print(f"frag {f} nmo {TA.shape[1]}: B finite {np.isfinite(Bs[0]).all()} sym finite {np.isfinite(Ss[0]).all()}",
          "B rep diffs", [float(np.abs(Bs[0] - b).max()) for b in Bs[1:]],
          "sym rep diffs", [float(np.abs(Ss[0] - s_).max()) for s_ in Ss[1:]])
    # independent check of K1 on a few columns with numpy
    off = 0; worst = 0.0
    rng = np.random.default_rng(0)
    cols = []
    for i, a in enumerate(A):
        for mu in a:
            cols.append((i, mu))
    pick = rng.choice(len(cols), size=40, replace=False)
    from quemb_b200 import indexing
    Tcols = T.exch_reachable_with_offsets
    M = p.unique_dense_data
    for c in pick:
        i, mu = cols[c]
        ref = np.zeros(p.naux)
        for col, nu in Tcols[mu]:
            ref += TA[nu, i] * M[:, col]
        worst = max(worst, float(np.abs(ref - Bs[0][:, c]).max()))
    print("   K1 vs numpy on 40 columns: max abs", worst)
