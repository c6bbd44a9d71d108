"""ncu target: K3 (gram_kernel) alone at the C32H66-like fragment shape (npair 24310 = nmo 220, naux 2716).
Launches the kernel `iters + 1` times on a constant X; prints executed TF/s (tile-padded flops)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from quemb_b200 import eri_sparse_DF as q

n = int(sys.argv[1]) if len(sys.argv) > 1 else 24310
k = int(sys.argv[2]) if len(sys.argv) > 2 else 2716
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 2
ctx = q.default_context(0)
print(f"gram_kernel n={n} k={k}: {ctx.probe_dmma_tflops(n, k, iters):.2f} TF/s (executed, tile padded)")
