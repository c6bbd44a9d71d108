"""Ad-hoc GPU probe (development aid, not a bench): FP64 peaks + per-kernel times of a few fragments."""
import sys, time, json
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch
from quemb_b200 import eri_sparse_DF as q, synthetic as syn, _lib

def dgemm_peak(n=8192, iters=5):
    a = torch.randn(n, n, dtype=torch.float64, device="cuda"); b = torch.randn(n, n, dtype=torch.float64, device="cuda")
    c = a @ b; torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 0
    for _ in range(iters):
        e0.record(); c = a @ b; e1.record(); torch.cuda.synchronize()
        best = max(best, 2 * n**3 / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    return best

which = sys.argv[1] if len(sys.argv) > 1 else "c32"
nfrag = int(sys.argv[2]) if len(sys.argv) > 2 else 3
whiten = (sys.argv[3] == "1") if len(sys.argv) > 3 else False
print("cuBLAS DGEMM 8192^3 TF/s:", dgemm_peak())
ctx = q.default_context(0)
print("DMMA probe TF/s (n=4096,k=2048):", ctx.probe_dmma_tflops(4096, 2048, 5))
print("DMMA probe TF/s (n=8192,k=2720):", ctx.probe_dmma_tflops(8192, 2720, 3))
t = time.time()
p = syn.c32h66_like(n_frag=nfrag + 1) if which == "c32" else syn.octane_like()
print("gen", time.time() - t, "nao", p.nao, "naux", p.naux, "n_unique", p.n_unique)
t = time.time()
T = q.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable)
L = q.GPU_MatrixHandle(p.L_PQ)
print("upload", time.time() - t)
if whiten:
    t = time.time(); T.whiten(L); print("whiten s", time.time() - t, "TF/s", p.naux**2 * p.n_unique / (time.time() - t) / 1e12)
ctx.set_timing(True)
ws = q.default_workspace(ctx)
for f, TA in enumerate(p.TAs[1:1 + nfrag] if which == "c32" else p.TAs[:nfrag]):
    nmo = TA.shape[1]; npair = nmo * (nmo + 1) // 2
    out = torch.empty((npair, npair), dtype=torch.float64, device="cuda")
    for rep in range(2):
        t = time.time()
        _lib.check(_lib.lib().qemb_fragment_transform(ws._h, T._h, None if whiten else L._h, q._screen_for(p.S_abs, ctx)._h,
            _lib.ptr(TA), TA.shape[0], TA.shape[0], nmo, 1e-5, 0, -1, _lib.c_p(out.data_ptr()), npair, 1, None))
        ctx.synchronize(); wall = time.time() - t
    ms = ws.times_ms(); wk = ws.work()
    print(f"frag {f} nmo {nmo} npair {npair} wall {wall*1e3:.1f} ms  times {ms}")
    for k, fk in (("K1", "F_K1"), ("K2a", "F_K2a"), ("K2b", "F_K2b"), ("K3", "F_K3")):
        if ms[k] > 0:
            print(f"   {k}: {wk[fk]/ (ms[k]*1e-3)/1e12:.2f} TF/s  ({wk[fk]:.3e} flop, {wk['B_'+k[:3]] if 'B_'+k[:3] in wk else 0:.3e} B)")
    Gs = out[:4, :4].cpu().numpy(); print("   G[:2,:2]", Gs[:2, :2].ravel(), "sym", bool(torch.equal(out, out.T)))
