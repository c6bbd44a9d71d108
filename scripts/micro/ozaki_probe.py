"""Feasibility probe for DESIGN section 8, item 1: K3 (G = X^T X, FP64) through int8 tensor-core GEMMs.

X (naux x npair) is scaled per column by a power of two and split into S slices of 7-bit integers (error-free:
X = s_j * sum_a 2^-(7a+6) D_a + remainder below 2^-(7S-1)).  Then
    G_ij = s_i s_j * sum_{d < S} 2^-(7d+12) * sum_{a+b=d} (D_a^T D_b)_ij ,
every inner sum exact in int32 (|D| <= 64, naux * 64^2 * S < 2^31), so the only errors are the dropped slices.
The probe uses torch._int_mm (cuBLASLt) for the integer GEMMs and plain torch for slicing / recombination: it
measures (a) the deviation from the FP64 product on K3-shaped data and (b) the time of the integer GEMMs alone.
Nothing in the product uses this; a product version needs its own tcgen05 kind::i8 kernel with the slicing fused
into the K2b epilogue and the recombination into the K3 epilogue.

    python scripts/micro/ozaki_probe.py [npair] [naux]
"""
import sys
import time

import torch

npair = int(sys.argv[1]) if len(sys.argv) > 1 else 11936
naux = int(sys.argv[2]) if len(sys.argv) > 2 else 2720
dev = "cuda"
torch.manual_seed(7)
# K3-like operand: X = L^-1 sym; columns (orbital pairs) of very different norms, rows (aux functions) decaying
col = torch.exp(-6.0 * torch.rand(npair, dtype=torch.float64, device=dev))
row = torch.exp(-4.0 * torch.rand(naux, dtype=torch.float64, device=dev))
X = torch.randn(naux, npair, dtype=torch.float64, device=dev) * col[None, :] * row[:, None]

ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
ev[0].record()
G_ref = X.T @ X
ev[1].record()
torch.cuda.synchronize()
t_f64 = ev[0].elapsed_time(ev[1]) * 1e-3
gmax = G_ref.abs().max().item()
print(f"npair {npair} naux {naux}: FP64 GEMM (full, cuBLAS) {t_f64 * 1e3:.2f} ms = {2.0 * npair * npair * naux / t_f64 / 1e12:.1f} TF/s; max|G| {gmax:.3e}")

for S in (5, 6, 7, 8):
    t0 = time.time()
    s = torch.exp2(torch.ceil(torch.log2(X.abs().amax(dim=0).clamp_min(1e-300))))  # per column, power of two
    R = X / s[None, :]                                                               # |R| <= 1, exact
    D = []
    scale = 64.0
    for a in range(S):
        Da = torch.trunc(R * scale)
        R = R * scale - Da                                                           # exact in FP64
        D.append(Da.to(torch.int8))
        scale = 128.0
    del R
    torch.cuda.synchronize()
    t_slice = time.time() - t0
    # integer GEMMs: for each d one product with the slices concatenated along K (exact int32 accumulation)
    Dt = [d.T.contiguous() for d in D]                                               # (npair, naux)
    acc = torch.zeros(npair, npair, dtype=torch.float64, device=dev)
    t_int = 0.0
    kpad = (-naux) % 8
    for d in range(S):
        A = torch.cat([Dt[a] for a in range(d + 1)], dim=1)                          # (npair, (d+1) naux)
        B = torch.cat([D[d - a] for a in range(d + 1)], dim=0)                       # ((d+1) naux, npair)
        if A.shape[1] % 8:
            pad = (-A.shape[1]) % 8
            A = torch.nn.functional.pad(A, (0, pad))
            B = torch.nn.functional.pad(B, (0, 0, 0, pad))
        torch._int_mm(A, B)
        torch.cuda.synchronize()
        ev[0].record()
        P = torch._int_mm(A, B)
        ev[1].record()
        torch.cuda.synchronize()
        t_int += ev[0].elapsed_time(ev[1]) * 1e-3
        acc += P.to(torch.float64) * (2.0 ** -(7 * d + 12))
        del A, B, P
    G = acc * s[:, None] * s[None, :]
    err = (G - G_ref).abs().max().item()
    print(f"S={S}: max |G_int8 - G_f64| = {err:.3e} ({err / gmax:.2e} of max|G|); int8 GEMMs (full matrices) {t_int * 1e3:.2f} ms "
          f"-> lower triangle only ~{t_int * 0.5 * 1e3:.2f} ms vs FP64 lower triangle ~{t_f64 * 0.5 * 1e3:.2f} ms; slicing in torch {t_slice * 1e3:.0f} ms")
    del D, Dt, acc, G
    torch.cuda.empty_cache()
