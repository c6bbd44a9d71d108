"""oracle/make_golden.py -- generates tests/golden/*.npz from the REFERENCE ITSELF.

Run in the build container (needs /root/reference -> oracle/_ref, see oracle/Makefile):

    python oracle/make_golden.py

For a handful of small seeded cases it stores the inputs and what the reference's own compiled
extension (quemb.molbe._cpp.eri_sparse_DF built unmodified from /root/reference) returns for every
step of the hot path: index maps (exch_reachable_unique, offsets), AO_by_MO, K1 (contract_with_TA_1st),
K2a (contract_with_TA_2nd_to_sym_dense) and the whole transform_integral.  These fixtures pin the
C restatement (tests/test_oracle.py, CPU) and the CUDA path (tests/test_gpu_parity.py, GPU); the GPU
box has no /root/reference, so the vectors travel instead.

Also stores the index golden of the reference's tests/test_indexing.py (ravel/unravel round trip).
"""

from __future__ import annotations

import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

from oracle import oracle  # noqa: E402
from quemb_b200 import synthetic as syn  # noqa: E402

OUT = ROOT / "tests" / "golden"


def csr(lists):
    return oracle.to_csr(lists)


def run_case(ref, name: str, p: syn.Problem, exch=None) -> None:
    exch = [np.asarray(x, dtype=np.int64) for x in (p.exch_reachable if exch is None else exch)]
    T = ref.SemiSparseSym3DTensor(
        np.asfortranarray(p.unique_dense_data), (p.naux, p.nao, p.nao), [x.tolist() for x in exch]
    )
    d: dict[str, np.ndarray] = {}
    d["nao"], d["naux"] = np.int64(p.nao), np.int64(p.naux)
    d["S_abs"], d["L_PQ"], d["M"] = p.S_abs, p.L_PQ, p.unique_dense_data
    d["eps"] = np.float64(p.MO_coeff_epsilon)
    d["exch_ptr"], d["exch_idx"] = csr(exch)
    d["uniq_ptr"], d["uniq_idx"] = csr(T.exch_reachable_unique)
    off = T.offsets
    keys = np.array(sorted(off.keys()), dtype=np.int64)
    d["offset_keys"], d["offset_vals"] = keys, np.array([off[k] for k in keys], dtype=np.int64)
    d["n_frag"] = np.int64(len(p.TAs))
    for f, TA in enumerate(p.TAs):
        A = ref.get_AO_per_MO(TA, p.S_abs, p.MO_coeff_epsilon)
        B = ref.contract_with_TA_1st(TA, T, A)
        sym = ref.contract_with_TA_2nd_to_sym_dense(B, TA)
        G = ref.transform_integral(T, TA, p.S_abs, p.L_PQ, p.MO_coeff_epsilon)
        d[f"TA{f}"] = TA
        d[f"A_ptr{f}"], d[f"A_idx{f}"] = csr(A)
        d[f"B{f}"] = np.asfortranarray(B.dense_data)
        d[f"sym{f}"] = np.asfortranarray(sym)
        d[f"G{f}"] = np.asfortranarray(G)
        print(f"  {name} frag {f}: nmo={TA.shape[1]} npair={G.shape[0]} nB={B.dense_data.shape[1]} "
              f"max|G|={np.abs(G).max():.3f}")
    np.savez_compressed(OUT / f"{name}.npz", **d)


def main() -> None:
    ref = oracle.reference_module()
    if ref is None:
        raise SystemExit("oracle/_ref not built: run `make -C oracle` in the build container")
    OUT.mkdir(parents=True, exist_ok=True)

    # 1. sparse chain (alkane-like), two fragments (edge + interior)
    p = syn.make_chain_problem(n_units=6, ao_per_unit=5, n_end=1, naux=40, frag_halfwidth=1,
                               n_bath=4, sigma=0.45, lam=0.7, cond=1e8, seed=11, name="chain_small")
    p.TAs = [p.TAs[0], p.TAs[3]]
    run_case(ref, "chain_small", p)

    # 2. 3-D lattice (water-cluster-like)
    p = syn.make_lattice_problem(n_side=3, ao_per_site=2, naux=48, n_frag=2, n_bath=5, sigma=0.45,
                                 lam=0.6, cond=1e6, seed=12)
    run_case(ref, "lattice_small", p)

    # 3. coarse MO screening (eps = 1e-2): makes the A(min)/TA(max) convention of K2a
    #    (eri_sparse_DF.cpp:595 + indexers.hpp:95) visible at O(1e-2)
    p = syn.make_chain_problem(n_units=5, ao_per_unit=4, n_end=0, naux=24, frag_halfwidth=1,
                               n_bath=6, sigma=0.6, lam=1.2, cond=1e4, seed=13,
                               MO_coeff_epsilon=1e-2, name="coarse_eps")
    p.TAs = [p.TAs[2]]
    run_case(ref, "coarse_eps", p)

    # 4. asymmetric reachability (the on-fly variant: only AOs near the fragment are sources,
    #    eri_sparse_DF.py:234-240, 436): missing keys read column 0 (cpp:271)
    p = syn.make_chain_problem(n_units=5, ao_per_unit=4, n_end=0, naux=24, frag_halfwidth=0,
                               n_bath=3, sigma=0.45, lam=0.5, cond=1e4, seed=14, name="asym_sources")
    TA = p.TAs[1]
    p.TAs = [TA]
    sources = np.nonzero((p.S_abs @ np.abs(TA)).max(axis=1) > 1e-3)[0]
    exch = [p.exch_reachable[mu] if mu in set(sources.tolist()) else np.zeros(0, np.int64)
            for mu in range(p.nao)]
    # rebuild M = L @ Y for the reduced unique set
    rng = np.random.default_rng(141)
    mus, _ = syn.unique_pairs(exch)
    p.unique_dense_data = np.asfortranarray(
        np.tril(p.L_PQ) @ (rng.standard_normal((p.naux, mus.size)) / np.sqrt(p.naux)))
    run_case(ref, "asym_sources", p, exch=exch)

    # 5. the reference's own index golden (tests/test_indexing.py:4-11), n_orbitals = 10
    rows = []
    n = 10
    for a in range(n):
        for b in range(a + 1):
            for c in range(a + 1):
                for d in range(c + 1 if a > c else b + 1):
                    rows.append((a, b, c, d))
    np.savez_compressed(OUT / "indexing_n10.npz", abcd=np.array(rows, dtype=np.int64))
    print("wrote", sorted(x.name for x in OUT.glob("*.npz")))


if __name__ == "__main__":
    main()
