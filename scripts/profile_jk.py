"""ncu / timing target: the J/K consumer kernels on a C32H66-like fragment (nmo 201, npair 20301, G = 3.3 GB).

    python scripts/profile_jk.py [stream|rows|both]      (QEMB_JK_KERNEL picks jk_stream.cu or jk.cu)
"""
import os
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
from quemb_b200 import eri_sparse_DF as q, synthetic as syn

which = sys.argv[1] if len(sys.argv) > 1 else "both"
p = syn.c32h66_like(n_frag=2)
T = q.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable)
T.whiten(p.L_PQ)
TA = p.TAs[1]
n = TA.shape[1]
ws = q.transform_integral_resident(T, TA, p.S_abs, None, p.MO_coeff_epsilon)
D = np.random.default_rng(0).standard_normal((n, n)); D = D + D.T
res = {}
for kern in (["stream", "rows"] if which == "both" else [which]):
    os.environ["QEMB_JK_KERNEL"] = kern
    for _ in range(4):
        vj, vk = ws.jk(D)
        ms, nb = ws.jk_time()
        print(f"jk[{kern}] nmo {n}: {ms:.3f} ms, {nb / ms / 1e6:.0f} GB/s of G", flush=True)
    res[kern] = (vj, vk)
if len(res) == 2:
    dj = np.abs(res["stream"][0] - res["rows"][0]).max() / np.abs(res["rows"][0]).max()
    dk = np.abs(res["stream"][1] - res["rows"][1]).max() / np.abs(res["rows"][1]).max()
    print(f"stream vs rows: rel dJ {dj:.2e} rel dK {dk:.2e}")
