"""Development aid: where the wall time of one end-to-end build goes (host view).
usage: python scripts/e2e_timeline.py [triangle=1] [n_host_buffers=3] [nfrag=32]"""
import sys, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import bench
from quemb_b200 import eri_sparse_DF as q
from quemb_b200.driver import FragmentERIBuilder

tri = int(sys.argv[1]) if len(sys.argv) > 1 else 1
nb = int(sys.argv[2]) if len(sys.argv) > 2 else 3
nfrag = int(sys.argv[3]) if len(sys.argv) > 3 else 32
frac = float(sys.argv[4]) if len(sys.argv) > 4 else 0.0
p = bench.make_workload("c32h66", use_gpu_matmul=True)
idx = list(range(len(p.TAs)))[:nfrag]
b = FragmentERIBuilder(0, n_host_buffers=nb)
ctx = b.ctx
ctx.set_output_mode(bool(tri))
ctx.set_full_column_fraction(frac)
T = q.SemiSparseSym3DTensor((p.naux, p.nao, p.nao), p.exch_reachable, ctx=ctx)
b.reserve_host(max(t.shape[1] for t in p.TAs))
log = []
orig_t, orig_w = q.transform_integral, b.ws.wait_output
def timed_t(*a, **k):
    t = time.perf_counter(); r = orig_t(*a, **k); log.append(("enqueue", time.perf_counter() - t)); return r
def timed_w(ticket):
    t = time.perf_counter(); r = orig_w(ticket); log.append(("wait", time.perf_counter() - t)); return r
q.transform_integral = timed_t; b.ws.wait_output = timed_w
chk = []
def sink(i, G):
    t = time.perf_counter(); chk.append(float(G[0, 0]) + float(G[-1, 0]) + float(G[0, -1])); log.append(("sink", time.perf_counter() - t))
for rep in range(3):
    log.clear()
    t0 = time.perf_counter()
    T.upload(p.unique_dense_data); ctx.synchronize(); t1 = time.perf_counter()
    b.stage(None, None, None, p.L_PQ, p.S_abs, whiten=True, tensor=T); ctx.synchronize(); t2 = time.perf_counter()
    b.build(p.TAs, p.MO_coeff_epsilon, sink, indices=idx); ctx.synchronize(); t3 = time.perf_counter()
    tot = {}
    for k, v in log: tot[k] = tot.get(k, 0) + v
    d2h = sum(8 * (p.TAs[i].shape[1] * (p.TAs[i].shape[1] + 1) // 2) ** 2 for i in idx)
    print(f"rep {rep} tri={tri} nb={nb} full_frac={frac}: upload {t1-t0:.3f} stage+whiten {t2-t1:.3f} build {t3-t2:.3f} total {t3-t0:.3f} s | "
          + " ".join(f"{k} {v:.3f}" for k, v in tot.items()) + f" | G bytes {d2h/1e9:.1f} GB -> {d2h/(t3-t2)/1e9:.1f} GB/s effective")
# symmetry / parity spot check of the last fragment against the full-square mode
i = idx[-1]; nmo = p.TAs[i].shape[1]; npair = nmo * (nmo + 1) // 2
G1 = np.array(q.transform_integral(T, p.TAs[i], b.screen, None, p.MO_coeff_epsilon, ws=b.ws))
ctx.set_output_mode(False)
G0 = np.array(q.transform_integral(T, p.TAs[i], b.screen, None, p.MO_coeff_epsilon, ws=b.ws))
print("triangle-vs-full max abs diff", float(np.abs(G1 - G0).max()), "symmetric", bool(np.array_equal(G1, G1.T)))
