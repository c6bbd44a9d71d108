"""Where the wall time of the all-fragment consumer pipeline (FragmentERIBuilder.build_jk) goes, host view."""
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np

import bench
from quemb_b200 import consumer
from quemb_b200 import eri_sparse_DF as q
from quemb_b200.driver import FragmentERIBuilder

p = bench.make_workload("c32h66", use_gpu_matmul=True)
b = FragmentERIBuilder(0)
T = q.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable, ctx=b.ctx)
b.stage(None, None, None, p.L_PQ, p.S_abs, whiten=True, tensor=T)
rng = np.random.default_rng(11)
hf_dm = rng.standard_normal((p.nao, p.nao)); hf_dm = hf_dm + hf_dm.T
hf_veff = rng.standard_normal((p.nao, p.nao)); hf_veff = hf_veff + hf_veff.T
S = np.eye(p.nao)
rdms, nfr = {}, {}
for i, TA in enumerate(p.TAs):
    n = TA.shape[1]
    c = np.linalg.qr(rng.standard_normal((n, max(1, n // 3))))[0]
    rdms[i] = c @ c.T
    nfr[i] = n // 2
log = {}


def timed(name, fn):
    def f(*a, **k):
        t = time.perf_counter()
        r = fn(*a, **k)
        log[name] = log.get(name, 0.0) + time.perf_counter() - t
        return r
    return f


q.transform_integral_resident = timed("transform_enqueue_s", q.transform_integral_resident)
b.ws.jk_enqueue = timed("jk_enqueue_s", b.ws.jk_enqueue)
for rep in range(4):
    log.clear()
    b.ctx.synchronize()
    t0 = time.perf_counter()
    out = consumer.initialize_fragments(b, p.TAs, p.MO_coeff_epsilon, hf_dm, S, hf_veff, rdms, nfr)
    t1 = time.perf_counter()
    print(f"rep {rep}: initialize_fragments {t1 - t0:.3f} s | " + " ".join(f"{k} {v:.3f}" for k, v in log.items()), flush=True)
# the same with the densities prepared up front (no host BLAS inside the loop)
pre = {i: [0.5 * (rdms[i] + rdms[i].T), rdms[i]] for i in rdms}
for rep in range(2):
    log.clear()
    b.ctx.synchronize()
    t0 = time.perf_counter()
    b.build_jk(p.TAs, p.MO_coeff_epsilon, lambda i: pre[i])
    print(f"precomputed densities rep {rep}: build_jk {time.perf_counter() - t0:.3f} s | " + " ".join(f"{k} {v:.3f}" for k, v in log.items()), flush=True)
# kernels only, for scale
b.ctx.synchronize()
t0 = time.perf_counter()
b.build_resident(p.TAs, p.MO_coeff_epsilon)
b.ctx.synchronize()
print(f"build_resident (no consumers): {time.perf_counter() - t0:.3f} s")
