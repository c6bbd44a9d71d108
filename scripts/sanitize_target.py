"""compute-sanitizer target: every kernel of the path once on a reference golden case (tiny, so that racecheck /
synccheck / memcheck finish in minutes): K1, K2a, K2b, K3 (plain and split-K), whitening, block staging, J/K stream and
rows kernels, the opt-in int8-slice K3.  Results are checked against the reference goldens as usual."""
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
import numpy as np

from helpers import Golden
from quemb_b200 import eri_sparse_DF as q

which = sys.argv[1] if len(sys.argv) > 1 else "all"
g = Golden("chain_small")
ctx = q.default_context(0)
T = q.SemiSparseSym3DTensor(g.M, (g.naux, g.nao, g.nao), g.exch, ctx=ctx)
L = q.GPU_MatrixHandle(g.L, ctx=ctx)
ws = q.Workspace(ctx)
worst = 0.0
for f in range(min(g.n_frag, 2)):
    fr = g.frag(f)
    G = q.transform_integral(T, fr["TA"], g.S_abs, L, g.eps, ws=ws)
    worst = max(worst, float(np.abs(G - fr["G"]).max()))
    n = fr["TA"].shape[1]
    D = np.random.default_rng(f).standard_normal((n, n))
    D = D + D.T
    q.transform_integral_resident(T, fr["TA"], g.S_abs, L, g.eps, ws=ws)
    for kern in ("stream", "rows"):
        os.environ["QEMB_JK_KERNEL"] = kern
        ws.jk(D)
    os.environ.pop("QEMB_JK_KERNEL", None)
T.whiten(L)
fr = g.frag(0)
G = q.transform_integral(T, fr["TA"], g.S_abs, None, g.eps, ws=ws)
worst = max(worst, float(np.abs(G - fr["G"]).max()))
if which in ("all", "int8"):
    ctx.set_k3_int8_slices(7)
    G = q.transform_integral(T, fr["TA"], g.S_abs, None, g.eps, ws=ws)
    worst = max(worst, float(np.abs(G - fr["G"]).max()))
    ctx.set_k3_int8_slices(0)
T2 = q.SemiSparseSym3DTensor((g.naux, g.nao, g.nao), g.exch, ctx=ctx)
blk = np.random.default_rng(0).standard_normal((5, 7, g.naux))
T2.scatter_block(blk, 3, 0)
ctx.synchronize()
print(f"sanitize target done, worst |G - golden| = {worst:.3e}")
assert worst <= 1e-10
