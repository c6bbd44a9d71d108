"""Per-step parity of the B200 path against the compiled reference (oracle/_ref) at full octane / C32H66 sizes:
where (which step, which column) does the largest deviation sit?  Development aid."""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np

from oracle import oracle
from quemb_b200 import eri_sparse_DF as q
from quemb_b200 import synthetic as syn

ref = oracle.reference_module()
which = sys.argv[1] if len(sys.argv) > 1 else "octane"
p = syn.octane_like() if which == "octane" else syn.c32h66_like(n_frag=2)
print(which, "nao", p.nao, "naux", p.naux, "n_unique", p.n_unique, "nmo", [t.shape[1] for t in p.TAs])
T_ref = ref.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), [x.tolist() for x in p.exch_reachable])
T = q.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable)
L = q.GPU_MatrixHandle(p.L_PQ)
frags = [0, 3] if which == "octane" else [0]
for f in frags:
    TA = p.TAs[f]
    nmo = TA.shape[1]
    A = q.get_AO_per_MO(TA, p.S_abs, p.MO_coeff_epsilon)
    A_ref = [list(a) for a in ref.get_AO_per_MO(TA, p.S_abs, p.MO_coeff_epsilon)]
    print(f"frag {f} nmo {nmo}: lists equal {A == A_ref}, sum|A| {sum(map(len, A))}, |A(i)| min/max {min(map(len, A))}/{max(map(len, A))}")
    B = q.contract_with_TA_1st(TA, T, A)
    B_ref = ref.contract_with_TA_1st(TA, T_ref, A_ref)
    dB = np.abs(B.dense_data - B_ref.dense_data)
    cols = np.nonzero(dB.max(axis=0) > 1e-11)[0]
    print(f"  K1: max err {dB.max():.3e}; bad columns {cols.size} of {dB.shape[1]}; first {cols[:12].tolist()}; rows of worst {np.nonzero(dB[:, dB.max(axis=0).argmax()] > 1e-11)[0][:8].tolist()}")
    if cols.size:
        offs = np.cumsum([0] + [len(a) for a in A])
        for c in cols[:6]:
            i = int(np.searchsorted(offs, c, side="right") - 1)
            print(f"     col {c}: MO {i}, k {c - offs[i]} of {len(A[i])}, mu {A[i][c - offs[i]]}")
    sym = q.contract_with_TA_2nd_to_sym_dense(B, TA)
    sym_ref = ref.contract_with_TA_2nd_to_sym_dense(B_ref, TA)
    dS = np.abs(sym - sym_ref)
    cols = np.nonzero(dS.max(axis=0) > 1e-11)[0]
    print(f"  K2a: max err {dS.max():.3e}; bad columns {cols.size} of {dS.shape[1]}; first {cols[:12].tolist()}")
    for c in cols[:6]:
        hi = int((np.sqrt(8 * c + 1) - 1) / 2)
        print(f"     col {c}: (hi {hi}, lo {c - hi * (hi + 1) // 2})  rows bad {np.nonzero(dS[:, c] > 1e-11)[0][:6].tolist()}")
    G_ref = ref.transform_integral(T_ref, TA, p.S_abs, p.L_PQ, p.MO_coeff_epsilon)
    G = q.transform_integral(T, TA, p.S_abs, L, p.MO_coeff_epsilon)
    dG = np.abs(G - G_ref)
    bad = np.argwhere(dG > 1e-10)
    print(f"  G (per-fragment solve): max err {dG.max():.3e}; bad entries {len(bad)}; rows {np.unique(bad[:, 0])[:10].tolist()} cols {np.unique(bad[:, 1])[:10].tolist()}")
    X = np.linalg.solve(np.tril(p.L_PQ), sym_ref)
    G_np = X.T @ X
    print(f"  numpy X^T X from the reference's sym vs reference G: {np.abs(G_np - G_ref).max():.3e}; vs B200 G: {np.abs(G_np - G).max():.3e}")
Tw = q.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable)
Tw.whiten(L)
for f in frags:
    TA = p.TAs[f]
    G_ref = ref.transform_integral(T_ref, TA, p.S_abs, p.L_PQ, p.MO_coeff_epsilon)
    Gw = q.transform_integral(Tw, TA, p.S_abs, None, p.MO_coeff_epsilon)
    print(f"frag {f} whitened tensor: max err {np.abs(Gw - G_ref).max():.3e}")
