"""Where do the sporadic wrong (ij|kl) entries come from: K3 itself or the device->host path?  Development aid."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import ctypes

import numpy as np
import torch

from quemb_b200 import _lib
from quemb_b200 import eri_sparse_DF as q
from quemb_b200 import synthetic as syn
from quemb_b200._lib import c_i64, check, lib, ptr

which = sys.argv[1] if len(sys.argv) > 1 else "octane"
p = syn.octane_like() if which == "octane" else syn.c32h66_like(n_frag=2)
T = q.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable)
L = q.GPU_MatrixHandle(p.L_PQ)
ctx = q.default_context()
ws = q.default_workspace(ctx)


def describe(tag, d, thr=1e-10):
    bad = np.argwhere(d > thr)
    if len(bad) == 0:
        print(f"    {tag}: ok (max {d.max():.2e})")
        return
    r, c = bad[:, 0], bad[:, 1]
    print(f"    {tag}: {len(bad)} bad, max {d.max():.2e}, rows {r.min()}..{r.max()} ({len(np.unique(r))} distinct), "
          f"cols {c.min()}..{c.max()} ({len(np.unique(c))} distinct); tiles {sorted(set(zip((r // 128).tolist(), (c // 128).tolist())))[:8]}")


for f in ([0, 3] if which == "octane" else [0, 1]):
    TA = p.TAs[f]
    nmo = TA.shape[1]
    npair = nmo * (nmo + 1) // 2
    print(f"frag {f} nmo {nmo} npair {npair}")
    for rep in range(4):
        q.transform_integral_resident(T, TA, p.S_abs, L, p.MO_coeff_epsilon, ws=ws)
        X = np.empty((p.naux, npair), order="F")
        check(lib().qemb_fragment_get_sym(ws._h, ptr(X), c_i64(p.naux)))
        Gd = ws.G(npair)
        Xt = torch.from_numpy(np.ascontiguousarray(X)).cuda()
        Gt = (Xt.T @ Xt).cpu().numpy()
        print(f"  rep {rep}")
        describe("resident G vs torch X^T X", np.abs(Gd - Gt))
        Gh = q.transform_integral(T, TA, p.S_abs, L, p.MO_coeff_epsilon, ws=ws)
        describe("host-path G (pageable) vs torch", np.abs(Gh - Gt))
        Gp = _lib.pinned_empty((npair, npair))
        q.transform_integral(T, TA, p.S_abs, L, p.MO_coeff_epsilon, ws=ws, out=Gp)
        describe("host-path G (pinned) vs torch", np.abs(Gp - Gt))
        ctx.set_output_mode(False)
        Gf = q.transform_integral(T, TA, p.S_abs, L, p.MO_coeff_epsilon, ws=ws)
        ctx.set_output_mode(True)
        describe("host-path G (full square D2H) vs torch", np.abs(Gf - Gt))
        _lib.pinned_free(Gp)
