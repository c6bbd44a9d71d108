"""What an Ozaki-style K3 could run at: int8 x int8 -> int32 GEMM rate of the library path (torch._int_mm ->
cuBLASLt) on K3's shape (npair x npair x naux) next to the FP64 DGEMM rate, on this GPU.  Informational only
(DESIGN section 8, item 1): nothing in the product uses it.

    python scripts/micro/int8_gemm_rate.py
"""
import json
import torch

dev = "cuda"
res = {}


def timed(fn, iters=5):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e-3


for (m, n, k) in ((20304, 20304, 2720), (11936, 11936, 2720), (8192, 8192, 8192)):
    a = torch.randint(-127, 127, (m, k), dtype=torch.int8, device=dev)
    b = torch.randint(-127, 127, (k, n), dtype=torch.int8, device=dev)
    t = timed(lambda: torch._int_mm(a, b))
    res[f"int8 {m}x{n}x{k}"] = {"s": t, "Top/s": 2.0 * m * n * k / t / 1e12}
    del a, b
    x = torch.randn(k, m, dtype=torch.float64, device=dev)
    t = timed(lambda: x.T @ x, iters=2)
    res[f"f64 {m}x{n}x{k}"] = {"s": t, "TFLOP/s": 2.0 * m * n * k / t / 1e12}
    del x
    torch.cuda.empty_cache()
for k_, v in res.items():
    print(k_, json.dumps(v))
