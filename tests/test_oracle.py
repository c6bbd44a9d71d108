"""CPU: pins the plain-C restatement (oracle/eri_oracle.c) against what the reference's own compiled
extension produced (tests/golden/*.npz, made by oracle/make_golden.py), and -- when oracle/_ref is
present (build container) -- against the live reference on a fresh seeded case."""
from __future__ import annotations

import numpy as np
import pytest

from oracle import oracle
from quemb_b200 import indexing, synthetic as syn

from helpers import CASES, Golden

TOL = 1e-10  # max-abs on (ij|kl); BASELINE.json north_star


@pytest.mark.parametrize("name", CASES)
def test_index_maps_bit_exact(name):
    g = Golden(name)
    rptr, ridx, rcol, uptr, uidx = oracle.reachable_columns(g.exch)
    assert oracle.from_csr(uptr, uidx) == [u.tolist() for u in g.uniq]
    # every key the reference assigned has the same column here
    for mu, lst in enumerate(g.exch):
        for k, nu in enumerate(lst):
            key = oracle.ravel_symmetric(mu, int(nu))
            assert rcol[rptr[mu] + k] == g.offsets.get(key, 0)
    # python host twins agree as well
    uniq = indexing.extract_unique(g.exch)
    assert uniq == [u.tolist() for u in g.uniq]
    off = indexing.unique_offsets(uniq)
    assert {k: v for k, v in g.offsets.items() if k in off} == off


@pytest.mark.parametrize("name", CASES)
def test_steps_match_reference_golden(name):
    g = Golden(name)
    for f in range(g.n_frag):
        fr = g.frag(f)
        A = oracle.get_AO_per_MO(fr["TA"], g.S_abs, g.eps)
        assert [a.tolist() for a in A] == [a.tolist() for a in fr["A"]], "AO_by_MO pattern must be bit-exact"
        B = oracle.contract_with_TA_1st(fr["TA"], g.M, g.exch, A)
        assert B.shape == fr["B"].shape
        assert np.abs(B - fr["B"]).max() <= 1e-12
        sym = oracle.contract_with_TA_2nd_to_sym_dense(B, fr["TA"], A)
        assert np.abs(sym - fr["sym"]).max() <= 1e-12
        G = oracle.transform_integral(g.M, g.exch, fr["TA"], g.S_abs, g.L, g.eps)
        assert G.shape == fr["G"].shape
        assert np.abs(G - fr["G"]).max() <= TOL
        assert np.array_equal(G, G.T)


def test_k2a_uses_lists_of_min_and_TA_of_max():
    """eri_sparse_DF.cpp:595 + indexers.hpp:95: with coarse screening the other convention is off by O(eps)."""
    g = Golden("coarse_eps")
    fr = g.frag(0)
    TA, A, B = fr["TA"], fr["A"], fr["B"]
    nmo = TA.shape[1]
    aptr, _ = oracle.to_csr(A)
    wrong = np.zeros_like(fr["sym"])
    for hi in range(nmo):
        for lo in range(hi + 1):
            col = indexing.ravel_symmetric(hi, lo)
            for k, mu in enumerate(A[hi]):  # lists of MAX (wrong)
                wrong[:, col] += TA[mu, lo] * B[:, aptr[hi] + k]
    assert np.abs(wrong - fr["sym"]).max() > 1e-6


def test_reference_indexing_golden(golden_dir):
    """tests/test_indexing.py:4-11 of the reference: unravel(ravel(a,b,c,d)) round trip, n = 10."""
    abcd = np.load(golden_dir / "indexing_n10.npz")["abcd"]
    for a, b, c, d in abcd.tolist():
        assert indexing.unravel_eri_idx(indexing.ravel_eri_idx(a, b, c, d)) == (a, b, c, d)
    for i in range(5000):
        assert oracle.unravel_symmetric(i) == indexing.unravel_symmetric(i)
        lo, hi = indexing.unravel_symmetric(i)
        assert lo <= hi and indexing.ravel_symmetric(lo, hi) == i == oracle.ravel_symmetric(hi, lo)


def test_dense_restatement_unscreened():
    """With complete lists the path equals the dense einsum / solve_triangular restatement."""
    from scipy.linalg import solve_triangular

    p = syn.make_chain_problem(n_units=3, ao_per_unit=3, n_end=0, naux=12, frag_halfwidth=0, n_bath=3,
                               sigma=5.0, cond=1e3, seed=5)
    nao = p.nao
    assert all(len(x) == nao for x in p.exch_reachable)
    mus, nus = syn.unique_pairs(p.exch_reachable)
    dense = np.zeros((p.naux, nao, nao))
    dense[:, mus, nus] = p.unique_dense_data
    dense[:, nus, mus] = p.unique_dense_data
    TA = p.TAs[1]
    nmo = TA.shape[1]
    Pij = np.einsum("Pmn,mi,nj->Pij", dense, TA, TA)
    sym = np.zeros((p.naux, nmo * (nmo + 1) // 2))
    for i in range(nmo):
        for j in range(i + 1):
            sym[:, indexing.ravel_symmetric(i, j)] = Pij[:, i, j]
    X = solve_triangular(p.L_PQ, sym, lower=True)
    G = oracle.transform_integral(p.unique_dense_data, p.exch_reachable, TA, np.ones((nao, nao)), p.L_PQ, 1e-5)
    assert np.abs(G - X.T @ X).max() <= 1e-11


def test_live_reference_if_present():
    ref = oracle.reference_module()
    if ref is None:
        pytest.skip("oracle/_ref not built on this box")
    p = syn.make_chain_problem(n_units=7, ao_per_unit=4, n_end=1, naux=32, frag_halfwidth=1, n_bath=5,
                               sigma=0.5, seed=77)
    T = ref.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao),
                                  [x.tolist() for x in p.exch_reachable])
    for TA in p.TAs[:3]:
        Gref = ref.transform_integral(T, TA, p.S_abs, p.L_PQ, p.MO_coeff_epsilon)
        G = oracle.transform_integral(p.unique_dense_data, p.exch_reachable, TA, p.S_abs, p.L_PQ,
                                      p.MO_coeff_epsilon)
        assert np.abs(G - Gref).max() <= TOL


def test_consumer_oracle_identity():
    """The two restatements in oracle/consumer_oracle.py agree with each other through
    e2[i] = sum_j D_ij (2J - K)_ij (the identity quemb_b200.consumer relies on), on a reference-produced G."""
    from oracle import consumer_oracle as co

    g = Golden("chain_small")
    fr = g.frag(0)
    G = fr["G"]
    nmo = fr["TA"].shape[1]
    rng = np.random.default_rng(3)
    C = np.linalg.qr(rng.standard_normal((nmo, nmo)))[0][:, : max(1, nmo // 3)]
    D = C @ C.T
    vj, vk = co.dot_eri_dm(G, D)
    assert np.allclose(vj, vj.T, atol=1e-13) and np.allclose(vk, vk.T, atol=1e-13)
    n_frag = max(1, nmo // 2)
    e2 = co.ebe_hf_e2(G, D, n_frag)
    assert np.abs(e2 - np.einsum("ij,ij->i", D[:n_frag], 2 * vj[:n_frag] - vk[:n_frag])).max() <= 1e-12


def test_consumer_oracle_two_independent_restatements_of_dot_eri_dm():
    """dot_eri_dm (PySCF, absent everywhere): the einsum restatement of its 4-index branch against explicit
    restore(1)-by-definition loops, on reference-produced (ij|kl) matrices, symmetric and non-symmetric dm."""
    from oracle import consumer_oracle as co

    rng = np.random.default_rng(11)
    for name, f in (("chain_small", 0), ("lattice_small", 1)):
        g = Golden(name)
        fr = g.frag(f)
        nmo_full = fr["TA"].shape[1]
        nmo = min(nmo_full, 10)
        # leading nmo x nmo block of orbitals: its pairs are the first nmo(nmo+1)/2 pair indices (T(max)+min order)
        npair = nmo * (nmo + 1) // 2
        G = np.ascontiguousarray(fr["G"][:npair, :npair])
        for sym in (True, False):
            dm = rng.standard_normal((nmo, nmo))
            if sym:
                dm = dm + dm.T
            vj, vk = co.dot_eri_dm(G, dm)
            wj, wk = co.dot_eri_dm_loops(G, dm)
            assert np.abs(vj - wj).max() <= 1e-12 * max(1.0, np.abs(wj).max())
            assert np.abs(vk - wk).max() <= 1e-12 * max(1.0, np.abs(wk).max())
        # restore(1) by loops == fancy-index unpacking, and carries the 8-fold symmetry of real ERIs
        e1, e2 = co.restore1_loops(G, nmo), co.unpack_s4(G, nmo)
        assert np.array_equal(e1, e2)
        assert np.array_equal(e1, e1.transpose(1, 0, 2, 3)) and np.array_equal(e1, e1.transpose(0, 1, 3, 2))
        assert np.abs(e1 - e1.transpose(2, 3, 0, 1)).max() <= 1e-14 * np.abs(e1).max()
        # J/K energy identities that need no unpacking convention: E_J = D:vj(D) = sum_{pq} Dp_p G_pq Dp_q with the
        # packed density (off-diagonals doubled)
        D = rng.standard_normal((nmo, nmo))
        D = D + D.T
        vj, vk = co.dot_eri_dm(G, D)
        tri = np.tril_indices(nmo)
        Dp = np.where(tri[0] == tri[1], 1.0, 2.0) * D[tri]
        assert abs(np.sum(D * vj) - Dp @ G @ Dp) <= 1e-11 * max(1.0, abs(Dp @ G @ Dp))


def test_k3_int8_bound_of_the_library_is_the_restatements():
    """The worst-case bound the guaranteed-tolerance mode decides with (``qemb_k3_int8_bound``) == the restatement's
    ``truncation_bound``, term by term; and it orders the slice counts."""
    from oracle import k3_int8_oracle as ko
    from quemb_b200 import eri_sparse_DF as q

    for naux in (1, 88, 700, 2716, 9024):
        prev = None
        for S in (5, 6, 7, 8):
            b = q.k3_int8_bound(naux, S)
            assert b == ko.truncation_bound(naux, S)
            assert prev is None or b < prev / 100
            prev = b


def _int8_test_operand(rng, naux=300, npair=90):
    X = rng.standard_normal((naux, npair)) * np.exp(-6 * rng.random(npair))[None, :] * np.exp(-4 * rng.random(naux))[:, None]
    X[:, 7] = 0.0                                   # an all-zero pair
    X[3, 11] = 0.75 * 2.0 ** 10                     # exactly representable extremes
    X[:, 13] = -np.abs(X[:, 13])                    # a one-signed column
    X[:, 17] *= 1e-300                              # scaled into the subnormal range (some entries underflow)
    X[5:40, 19] = (np.arange(35) - 17 + 0.5) * 2.0 ** -46  # exact ties of the 6-slice rounding (column maximum 0.75)
    X[0, 19], X[1:5, 19], X[40:, 19] = 0.75, 0.0, 0.0
    X[:, 23] = 1.0 - 2.0 ** -53                     # largest r below 1: N reaches 2^(8S-2)
    return X


def test_k3_int8_oracle_split_is_error_free_and_within_its_bound():
    """The restatement of the opt-in int8-slice K3: the split reconstructs X exactly (slices + remainder), the
    integer sums stay inside int32, and the recombined matrix is within the worst-case truncation bound of the FP64
    product (and far inside it in practice)."""
    from oracle import k3_int8_oracle as ko

    rng = np.random.default_rng(5)
    naux, npair = 300, 90
    X = _int8_test_operand(rng, naux, npair)
    G = X.T @ X
    for S in (5, 6, 7, 8):
        e, D, R = ko.slice_int8(X, S)
        N = np.zeros(X.shape, dtype=np.int64)
        for a in range(S):                           # the digits are those of one integer, N = rint(r 2^(8S-2)) ...
            N = N * 256 + D[a].astype(np.int64)
        rec = np.ldexp(N.astype(np.float64), -(8 * S - 2)) + R   # ... and N 2^-(8S-2) + R gives r back exactly
        assert np.array_equal(np.ldexp(rec, e[None, :]), X)
        assert all(d.min() >= -128 and d.max() <= 127 and np.array_equal(d, np.trunc(d)) for d in D)
        assert np.abs(D[0]).max() <= 65
        assert np.abs(R).max() <= 2.0 ** -(8 * S - 1)
        G8 = ko.gram_int8(X, S)
        assert np.array_equal(G8, G8.T)
        scale = np.ldexp(1.0, e[:, None] + e[None, :])
        ok = scale > 0                               # 2^(e_i+e_j) underflows for the subnormal column: G is 0 there
        assert np.array_equal(G8[~ok], G[~ok])
        assert (np.abs(G8 - G)[ok] / scale[ok]).max() <= ko.truncation_bound(naux, S) + 4 * naux * 2.0 ** -53
        assert np.abs(G8 - G).max() <= {5: 1e-9, 6: 1e-11, 7: 1e-14, 8: 1e-14}[S] * np.abs(G).max()
    # a NaN or an Inf in X poisons row and column of that pair, as the FP64 product would, and nothing else
    Xb = X.copy()
    Xb[5, 20], Xb[9, 33] = np.nan, np.inf
    Gb, G0 = ko.gram_int8(Xb, 7), ko.gram_int8(X, 7)
    keep = np.ones(npair, bool)
    keep[[20, 33]] = False
    assert np.isnan(Gb[~keep, :]).all() and np.isnan(Gb[:, ~keep]).all()
    assert np.array_equal(Gb[np.ix_(keep, keep)], G0[np.ix_(keep, keep)])
    # worst case (every rounding error aligned) at the benchmark's naux, per unit of 2^(e_i+e_j)
    assert 1e-9 > ko.truncation_bound(2716, 6) > 1e-10 > 2e-12 > ko.truncation_bound(2716, 7) and ko.truncation_bound(2716, 8) < 1e-14


def test_k3_int8_device_slicer_source_matches_the_restatement():
    """qemb_k3_int8_slice_host runs the digit routine the device slicer runs (csrc/k3_int8_digits.h, one source for
    host and device) under the same exponent rule: its exponents and digits must equal the restatement's bit for bit,
    including zero columns, subnormals, exact ties, N = 2^(8S-2) and NaN / Inf columns."""
    import ctypes

    from oracle import k3_int8_oracle as ko
    from quemb_b200._lib import lib

    rng = np.random.default_rng(6)
    X = np.asfortranarray(_int8_test_operand(rng))
    X[5, 20], X[9, 33] = np.nan, -np.inf
    naux, npair = X.shape
    fn = lib().qemb_k3_int8_slice_host
    for S in (5, 6, 7, 8):
        D = np.zeros((npair, S, naux), dtype=np.int8)
        expo = np.zeros(npair, dtype=np.int32)
        rc = fn(X.ctypes.data_as(ctypes.c_void_p), ctypes.c_int64(naux), ctypes.c_int64(naux), ctypes.c_int64(npair),
                ctypes.c_int(S), D.ctypes.data_as(ctypes.c_void_p), expo.ctypes.data_as(ctypes.c_void_p))
        assert rc == 0
        e, Dr, _ = ko.slice_int8(X, S)
        bad = ~np.isfinite(X).all(axis=0)
        assert np.array_equal(expo[~bad], e[~bad]) and (expo[bad] == ko.EXPO_POISON).all() and bad.sum() == 2
        for a in range(S):
            assert np.array_equal(D[:, a, :].T.astype(np.float64), Dr[a]), (S, a)


def test_k3_int8_parity_study_quick():
    """Reduced run of tests/int8_parity_study.py (the full sweep is profiles/k3_int8_parity_study_r1.jsonl): on real
    fragment operands, for metric condition numbers 1e2 .. 1e10 and for X cut into row slabs, 6 slices stay two
    orders of magnitude inside the 1e-10 bar, 7 slices at the FP64 product's own rounding level, and the 8-slice
    result agrees with long double to 1e-15."""
    import int8_parity_study as st

    rows = st.run(quick=True, out=None)
    assert len(rows) == 6
    for r in rows:
        assert r["exact_vs_longdouble"] <= 1e-15
        for ns in (1, 4):
            assert r[f"int8x6_slabs{ns}"] <= 1e-12 and r[f"int8x6_slabs{ns}"] <= r["bound_x_scale_6"]
            assert r[f"int8x7_slabs{ns}"] <= max(1e-15, 4.0 * r["fp64_vs_exact"])
