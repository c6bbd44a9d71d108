"""Stress check for the mbarrier pipelines (K3 and the K1/K2a gather GEMM): many repetitions of the same
fragment; K3 is compared entry by entry with a cuBLAS FP64 X^T X of the library's own X, K1/K2a bitwise between
repetitions (the kernels are deterministic) and once against the compiled reference.  Prints one summary line
per case; exits 1 on any deviation.  (Found the premature stage release fixed in round 1: ~1 row of 1 warp wrong
in 1 of 10^4 tiles.)"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch

from oracle import oracle
from quemb_b200 import eri_sparse_DF as q
from quemb_b200 import synthetic as syn
from quemb_b200._lib import c_i64, check, lib, ptr

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 25
ref = oracle.reference_module()
fail = 0
for which in ("octane", "c32"):
    p = syn.octane_like() if which == "octane" else syn.c32h66_like(n_frag=3)
    ctx = q.default_context()
    ws = q.default_workspace(ctx)
    T = q.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable)
    L = q.GPU_MatrixHandle(p.L_PQ)
    T_ref = None
    if ref is not None:
        T_ref = ref.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), [x.tolist() for x in p.exch_reachable])
    for f in ([0, 3] if which == "octane" else [0, 2]):
        TA = p.TAs[f]
        nmo = TA.shape[1]
        npair = nmo * (nmo + 1) // 2
        A = q.get_AO_per_MO(TA, p.S_abs, p.MO_coeff_epsilon)
        # K1 / K2a: bitwise reproducibility + reference
        B0 = S0 = None
        gdiff = 0
        for rep in range(reps):
            B = q.contract_with_TA_1st(TA, T, A, ws=ws)
            Bd = B.dense_data
            Sd = q.contract_with_TA_2nd_to_sym_dense(B, TA)
            if B0 is None:
                B0, S0 = Bd, Sd
            else:
                gdiff += int((Bd != B0).sum()) + int((Sd != S0).sum())
        eB = eS = float("nan")
        if T_ref is not None:
            B_ref = ref.contract_with_TA_1st(TA, T_ref, [list(a) for a in A])
            eB = float(np.abs(B0 - B_ref.dense_data).max())
            eS = float(np.abs(S0 - ref.contract_with_TA_2nd_to_sym_dense(B_ref, TA)).max())
        # K3
        bad_runs = bad_entries = 0
        worst = 0.0
        Gt = None
        for rep in range(reps):
            q.transform_integral_resident(T, TA, p.S_abs, L, p.MO_coeff_epsilon, ws=ws)
            if Gt is None:
                X = np.empty((p.naux, npair), order="F")
                check(lib().qemb_fragment_get_sym(ws._h, ptr(X), c_i64(p.naux)))
                Xt = torch.from_numpy(np.ascontiguousarray(X)).cuda()
                Gt = Xt.T @ Xt
            dev = ctypes_ptr = None
            import ctypes
            dptr, n_ = ctypes.c_void_p(), c_i64()
            check(lib().qemb_fragment_G_dev(ws._h, ctypes.byref(dptr), ctypes.byref(n_)))
            ctx.synchronize()
            Gd = torch.from_numpy(ws.G(npair)).cuda()
            d = (Gd - Gt).abs()
            nb = int((d > 1e-10).sum())
            worst = max(worst, float(d.max()))
            bad_runs += nb > 0
            bad_entries += nb
        ok = gdiff == 0 and bad_runs == 0 and (not (eB > 1e-12)) and (not (eS > 1e-12))
        fail += not ok
        print(f"{which} frag {f} nmo {nmo} npair {npair} reps {reps}: K1/K2a run-to-run differing entries {gdiff}, "
              f"K1 vs reference {eB:.2e}, K2a vs reference {eS:.2e}; K3 bad runs {bad_runs}/{reps}, bad entries {bad_entries}, "
              f"worst |G - X^T X| {worst:.2e}  {'OK' if ok else 'FAIL'}", flush=True)
sys.exit(1 if fail else 0)
