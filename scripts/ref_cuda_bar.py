"""The reference's OWN GPU variant as a bar (SURVEY 8d / acceptance 6): `transform_integral_cuda` of the unmodified
extension built with -DUSE_CUDA (oracle/_ref/cuda: CPU K1/K2a, then cublasDtrsm + cublasDsyrk with per-call
cudaMalloc, synchronous H2D/D2H and a serial host mirror, cpp:624-736) next to this repository's path on the same
fragments, same host buffers in and out.  Two builds of the reference are timed: "as shipped" (oracle/_ref/cuda: its
dsyrk adds to uninitialised memory, beta = 1, SURVEY F4 -- only its TIME is meaningful) and "beta fixed"
(oracle/_ref/cuda_betafix: the same sources with beta = 0, patched into a scratch copy by oracle/patch_beta.py), whose
OUTPUT is also compared with this repository's (ij|kl)."""
import importlib.util
import json
import sys
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import numpy as np

from quemb_b200 import eri_sparse_DF as q
from quemb_b200 import synthetic as syn

variant = sys.argv[1] if len(sys.argv) > 1 else "cuda"  # "cuda" (as shipped) or "cuda_betafix"
so = next((ROOT / "oracle" / "_ref" / variant).glob("eri_sparse_DF*.so"), None)
if so is None:
    raise SystemExit(f"oracle/_ref/{variant} not built")
spec = importlib.util.spec_from_file_location("eri_sparse_DF", str(so))
ref = importlib.util.module_from_spec(spec)
spec.loader.exec_module(ref)

p = syn.c32h66_like(n_frag=3)
T_ref = ref.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), [x.tolist() for x in p.exch_reachable])
L_ref = ref.GPU_MatrixHandle(p.L_PQ)
T = q.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable)
L = q.GPU_MatrixHandle(p.L_PQ)
Tw = q.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable)
Tw.whiten(L)
out = []
for f in (0, 2):
    TA = p.TAs[f]
    nmo = TA.shape[1]
    npair = nmo * (nmo + 1) // 2
    t_ref = []
    for _ in range(2):
        t0 = time.perf_counter()
        G_ref = ref.transform_integral_cuda(T_ref, TA, p.S_abs, L_ref, p.MO_coeff_epsilon)
        t_ref.append(time.perf_counter() - t0)
    t_b = []
    for _ in range(3):
        t0 = time.perf_counter()
        G = q.transform_integral(T, TA, p.S_abs, L, p.MO_coeff_epsilon)
        t_b.append(time.perf_counter() - t0)
    t_w = []
    for _ in range(3):
        t0 = time.perf_counter()
        G = q.transform_integral(Tw, TA, p.S_abs, None, p.MO_coeff_epsilon)
        t_w.append(time.perf_counter() - t0)
    parity = float(np.abs(G - G_ref).max()) if variant == "cuda_betafix" else None
    del G_ref
    row = {"fragment": f, "nmo": nmo, "npair": npair, "G_gb": npair * npair * 8 / 1e9, "reference_variant": variant,
           "max_abs_b200_vs_reference_gpu_variant": parity,
           "reference_transform_integral_cuda_s": min(t_ref),
           "b200_transform_integral_s": min(t_b), "b200_whitened_s": min(t_w),
           "speedup_vs_reference_gpu_variant": min(t_ref) / min(t_w),
           "note": "wall clock of one synchronous call, host arrays in, full (ij|kl) in pageable host memory out"}
    print(json.dumps(row), flush=True)
    out.append(row)
Path("gpurun_out").mkdir(exist_ok=True)
Path(f"gpurun_out/ref_cuda_bar_{variant}.json").write_text(json.dumps(out, indent=1))
