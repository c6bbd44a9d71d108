"""J/K consumer kernels over fragment sizes: device time of jk_stream.cu (default up to 256 orbitals) and jk.cu
(chunked rows, any size up to 768) on the (ij|kl) of a random nmo-orbital fragment; GB/s = 8 npair^2 / time.

    python scripts/jk_sweep.py [--out gpurun_out/jk_sweep.json]
"""
import argparse
import json
import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np

from quemb_b200 import eri_sparse_DF as q
from quemb_b200 import synthetic as syn

ap = argparse.ArgumentParser()
ap.add_argument("--out", default="gpurun_out/jk_sweep.json")
args = ap.parse_args()
p = syn.make_chain_problem(n_units=14, ao_per_unit=24, n_end=5, naux=96, frag_halfwidth=2, n_bath=60, sigma=0.94,
                           seed=311, n_frag=8)
T = q.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable)
T.whiten(p.L_PQ)
rng = np.random.default_rng(3)
rows = []
for nmo in (20, 50, 100, 150, 201, 225, 256, 300):
    TA = np.linalg.qr(rng.standard_normal((p.nao, nmo)))[0]
    ws = q.transform_integral_resident(T, TA, p.S_abs, None, 0.0)
    D = rng.standard_normal((nmo, nmo))
    D = D + D.T
    row = {"nmo": nmo, "npair": nmo * (nmo + 1) // 2}
    for kern in ("stream", "rows"):
        os.environ["QEMB_JK_KERNEL"] = kern
        best = None
        for _ in range(4):
            ws.jk(D)
            ms, nb = ws.jk_time()
            best = ms if best is None else min(best, ms)
        row[kern] = {"ms": best, "gbs": nb / best / 1e6}
    rows.append(row)
    print(json.dumps(row), flush=True)
Path(args.out).parent.mkdir(parents=True, exist_ok=True)
Path(args.out).write_text(json.dumps(rows, indent=1))
