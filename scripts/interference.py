"""Development aid: does the D2H copy slow down under concurrent K3 / cuBLAS DGEMM / host mirror threads?"""
import sys, time, threading, ctypes
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np, torch
from quemb_b200 import eri_sparse_DF as q, _lib

n = 24310
ctx = q.default_context(0)
G = torch.zeros(n * n, dtype=torch.float64, device="cuda")
H = torch.empty(n * n, dtype=torch.float64).pin_memory()
H2 = torch.zeros(n * n, dtype=torch.float64).pin_memory()
side = torch.cuda.Stream()
L = _lib.lib()

def d2h(reps=4):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(side):
        e0.record(side)
        for _ in range(reps):
            H.copy_(G, non_blocking=True)
        e1.record(side)
    e1.synchronize()
    return reps * n * n * 8 / (e0.elapsed_time(e1) * 1e-3) / 1e9

def bg(fn):
    t = threading.Thread(target=fn); t.start(); return t

def k3_load():
    print("   K3 TF/s under load:", round(ctx.probe_dmma_tflops(n, 2716, 12), 2))
def dgemm_load():
    a = torch.randn(8192, 8192, dtype=torch.float64, device="cuda"); b = a.clone()
    for _ in range(22): c = a @ b
    torch.cuda.synchronize()
def mirror_load():
    t = time.perf_counter()
    for _ in range(2):
        _lib.check(L.qemb_host_mirror_lower(ctypes.c_void_p(H2.data_ptr()), ctypes.c_int64(n), ctypes.c_int64(n)))
    dt = time.perf_counter() - t
    print(f"   mirror rate: {2 * n * n * 4 / dt / 1e9:.1f} GB/s mirrored")

print("D2H alone:", round(d2h(), 1), "GB/s")
print("K3 alone:"); k3_load()
print("mirror alone:"); mirror_load()
t = bg(k3_load); time.sleep(0.15); print("D2H under K3:", round(d2h(), 1), "GB/s"); t.join()
t = bg(dgemm_load); time.sleep(0.3); print("D2H under cuBLAS DGEMM:", round(d2h(), 1), "GB/s"); t.join()
t = bg(mirror_load); time.sleep(0.05); print("D2H under host mirror:", round(d2h(), 1), "GB/s"); t.join()
t = bg(k3_load); t2 = bg(mirror_load); time.sleep(0.15); print("D2H under K3 + mirror:", round(d2h(), 1), "GB/s"); t.join(); t2.join()
