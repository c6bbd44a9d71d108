"""GPU: parity of the sm_100a path (through the C ABI) with
  (1) the committed golden fixtures the reference's own compiled extension produced,
  (2) the plain-C oracle on fresh seeded inputs,
  (3) size-independent properties at larger sizes.
Tolerances: index maps / screening pattern bit-exact; (ij|kl) <= 1e-10 max-abs (BASELINE.json north_star)."""
from __future__ import annotations

import numpy as np
import pytest

from oracle import oracle
from quemb_b200 import _lib, indexing, synthetic as syn

from helpers import CASES, Golden

pytestmark = pytest.mark.gpu

TOL = 1e-10       # max-abs on (ij|kl)
TOL_STEP = 1e-12  # K1 / K2a intermediates (pure sums of O(1) products)


@pytest.fixture(scope="module")
def q():
    from quemb_b200 import eri_sparse_DF as mod

    mod.default_context(0)  # raises QembNoDevice without a B200: no CPU fallback
    return mod


def _tensor(q, g: Golden):
    return q.SemiSparseSym3DTensor(g.M, (g.naux, g.nao, g.nao), [x.tolist() for x in g.exch])


@pytest.mark.parametrize("name", CASES)
def test_index_maps_bit_exact(q, name):
    g = Golden(name)
    T = _tensor(q, g)
    assert T.exch_reachable_unique == [u.tolist() for u in g.uniq]
    off = T.offsets
    assert {k: v for k, v in g.offsets.items() if k in off} == off
    for mu, row in enumerate(T.exch_reachable_with_offsets):
        for col, nu in row:
            assert col == g.offsets.get(indexing.ravel_symmetric(mu, nu), 0)
    assert T.shape == (g.naux, g.nao, g.nao)
    assert T.nonzero_size == g.naux * g.M.shape[1]
    np.testing.assert_array_equal(T.unique_dense_data, g.M)
    mu = g.nao // 2
    nu = int(g.exch[mu][0]) if len(g.exch[mu]) else mu
    if len(g.exch[mu]):
        np.testing.assert_array_equal(T[mu, nu], g.M[:, g.offsets[indexing.ravel_symmetric(mu, nu)]])


@pytest.mark.parametrize("name", CASES)
def test_steps_match_reference_golden(q, name):
    g = Golden(name)
    T = _tensor(q, g)
    for f in range(g.n_frag):
        fr = g.frag(f)
        A = q.get_AO_per_MO(fr["TA"], g.S_abs, g.eps)
        assert A == [a.tolist() for a in fr["A"]], "AO_by_MO pattern must be bit-exact"
        B = q.contract_with_TA_1st(fr["TA"], T, A)
        assert B.dense_data.shape == fr["B"].shape
        assert np.abs(B.dense_data - fr["B"]).max() <= TOL_STEP
        sym = q.contract_with_TA_2nd_to_sym_dense(B, fr["TA"])
        assert np.abs(sym - fr["sym"]).max() <= TOL_STEP
        G = q.transform_integral(T, fr["TA"], g.S_abs, g.L, g.eps)
        assert G.shape == fr["G"].shape
        assert np.abs(G - fr["G"]).max() <= TOL
        assert np.array_equal(G, G.T)
        # GPU_MatrixHandle route (transform_integral_cuda of the reference)
        G2 = q.transform_integral_cuda(T, fr["TA"], g.S_abs, q.GPU_MatrixHandle(g.L), g.eps)
        assert np.array_equal(G, G2)


@pytest.mark.parametrize("name", CASES)
def test_K2a_on_caller_built_tensors_matches_reference_golden(q, name):
    """The reference's public SemiSparse3DTensor constructors (pyi:41-58) and mut_dense_data feed
    contract_with_TA_2nd_to_sym_dense from HOST data (qemb_fragment_set_B); expected values: the reference's own `sym`."""
    g = Golden(name)
    T = _tensor(q, g)
    for f in range(g.n_frag):
        fr = g.frag(f)
        A, TA, Bref = [a.tolist() for a in fr["A"]], fr["TA"], np.asfortranarray(fr["B"])
        shape = (g.naux, g.nao, TA.shape[1])
        # (dense_data, shape, lists)
        sym = q.contract_with_TA_2nd_to_sym_dense(q.SemiSparse3DTensor(Bref, shape, A), TA)
        assert np.abs(sym - fr["sym"]).max() <= TOL_STEP
        # (shape, lists) filled through mut_dense_data
        t = q.SemiSparse3DTensor(shape, A)
        assert np.isnan(t.dense_data).all() or t.nonzero_size == 0
        t.mut_dense_data[...] = Bref
        assert np.array_equal(q.contract_with_TA_2nd_to_sym_dense(t, TA), sym)
        # expert form with permuted columns
        n = Bref.shape[1]
        perm = np.random.default_rng(f).permutation(n)
        with_off, k = [], 0
        for row in A:
            with_off.append([(int(perm[k + j]), mu) for j, mu in enumerate(row)])
            k += len(row)
        D = np.empty_like(Bref)
        D[:, perm] = Bref
        e = q.SemiSparse3DTensor(D, shape, A, with_off, {mu + i * g.nao: o for i, r in enumerate(with_off) for o, mu in r})
        assert np.array_equal(q.contract_with_TA_2nd_to_sym_dense(e, TA), sym)
        # a K1 result: written through mut_dense_data (scaled by 2), and contracted with ANOTHER TA than K1 ran with
        B = q.contract_with_TA_1st(TA, T, A)
        assert B._resident()
        TA2 = np.asfortranarray(TA * 0.5)
        sym_half = q.contract_with_TA_2nd_to_sym_dense(B, TA2)  # same lists, other coefficients: host route
        assert np.abs(sym_half - 0.5 * fr["sym"]).max() <= TOL_STEP
        B = q.contract_with_TA_1st(TA, T, A)
        B.mut_dense_data[...] *= 2.0
        assert not B._resident()
        assert np.abs(q.contract_with_TA_2nd_to_sym_dense(B, TA) - 2.0 * fr["sym"]).max() <= 2 * TOL_STEP
        # a tensor whose workspace has moved on says so instead of returning another fragment's columns
        B = q.contract_with_TA_1st(TA, T, A)
        q.get_AO_per_MO(TA, g.S_abs, g.eps)
        with pytest.raises(_lib.QembError, match="reused"):
            B.dense_data


@pytest.mark.parametrize("name", CASES)
def test_whitened_tensor_matches_reference_golden(q, name):
    """Design option O1: L^-1 applied to (P|mu nu) once instead of per fragment."""
    g = Golden(name)
    T = _tensor(q, g)
    T.whiten(g.L)
    assert T.whitened
    for f in range(g.n_frag):
        fr = g.frag(f)
        G = q.transform_integral(T, fr["TA"], g.S_abs, None, g.eps)
        assert np.abs(G - fr["G"]).max() <= TOL
        assert np.array_equal(G, G.T)


def _oracle_G(p, TA):
    return oracle.transform_integral(p.unique_dense_data, p.exch_reachable, TA, p.S_abs, p.L_PQ, p.MO_coeff_epsilon)


@pytest.mark.parametrize(
    "kw",
    [
        dict(n_units=10, ao_per_unit=12, n_end=3, naux=333, frag_halfwidth=1, n_bath=20, sigma=0.7, cond=1e8, seed=101),
        dict(n_units=6, ao_per_unit=9, n_end=0, naux=129, frag_halfwidth=2, n_bath=9, sigma=0.5, cond=1e10, seed=102),
        dict(n_units=4, ao_per_unit=3, n_end=1, naux=17, frag_halfwidth=0, n_bath=2, sigma=0.4, cond=1e3, seed=103),
    ],
)
def test_seeded_cases_match_oracle(q, kw):
    """naux not a multiple of 16/128, several tiles in every dimension, ill-conditioned metric."""
    p = syn.make_chain_problem(**kw)
    T = q.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable)
    L = q.GPU_MatrixHandle(p.L_PQ)
    for TA in (p.TAs[0], p.TAs[len(p.TAs) // 2]):
        A_ref = oracle.get_AO_per_MO(TA, p.S_abs, p.MO_coeff_epsilon)
        assert q.get_AO_per_MO(TA, p.S_abs, p.MO_coeff_epsilon) == [a.tolist() for a in A_ref]
        G = q.transform_integral(T, TA, p.S_abs, L, p.MO_coeff_epsilon)
        Gref = _oracle_G(p, TA)
        assert np.abs(G - Gref).max() <= TOL
        assert np.array_equal(G, G.T)
    Tw = q.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable)
    Tw.whiten(L)
    G = q.transform_integral(Tw, p.TAs[0], p.S_abs, None, p.MO_coeff_epsilon)
    assert np.abs(G - _oracle_G(p, p.TAs[0])).max() <= TOL


def test_lattice_case_matches_oracle(q):
    p = syn.make_lattice_problem(n_side=4, ao_per_site=3, naux=200, n_frag=2, n_bath=12, sigma=0.5, seed=104)
    T = q.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable)
    for TA in p.TAs:
        G = q.transform_integral(T, TA, p.S_abs, p.L_PQ, p.MO_coeff_epsilon)
        assert np.abs(G - _oracle_G(p, TA)).max() <= TOL


def test_edge_cases(q):
    """single MO, MO that reaches no AO (empty list), AO with an empty reachable list."""
    rng = np.random.default_rng(7)
    nao, naux = 9, 21
    S = np.eye(nao)
    S[0, 1] = S[1, 0] = 0.5
    exch = [np.nonzero(S[:, mu] >= 1e-10)[0].astype(np.int64) for mu in range(nao)]
    exch[4] = np.zeros(0, np.int64)  # AO 4 reaches nothing
    mus, _ = syn.unique_pairs(exch)
    M = np.asfortranarray(rng.standard_normal((naux, mus.size)))
    L = np.asfortranarray(np.linalg.cholesky(np.eye(naux) * 2 + 0.1 * np.ones((naux, naux))))
    T = q.SemiSparseSym3DTensor(M, (naux, nao, nao), exch)
    # (a) one MO
    TA = np.zeros((nao, 1), order="F")
    TA[0, 0] = 1.0
    G = q.transform_integral(T, TA, S, L, 1e-5)
    assert G.shape == (1, 1)
    assert np.abs(G - oracle.transform_integral(M, exch, TA, S, L, 1e-5)).max() <= TOL
    # (b) second MO is identically zero -> empty A list; third sits on the AO with no partners
    TA = np.zeros((nao, 3), order="F")
    TA[1, 0] = 1.0
    TA[4, 2] = 1.0
    A = q.get_AO_per_MO(TA, S, 1e-5)
    assert A == [[0, 1], [], [4]]
    G = q.transform_integral(T, TA, S, L, 1e-5)
    Gref = oracle.transform_integral(M, exch, TA, S, L, 1e-5)
    assert np.abs(G - Gref).max() <= TOL
    # pair columns that involve the empty MO are exactly zero
    for hi, lo in [(1, 0), (1, 1), (2, 1)]:
        assert not G[indexing.ravel_symmetric(hi, lo)].any()


def test_error_behaviour(q):
    nao, naux = 6, 8
    exch = [list(range(nao))] * nao
    with pytest.raises(RuntimeError):
        q.SemiSparseSym3DTensor(np.zeros((naux, 3)), (naux, nao, nao), exch)  # wrong n_unique (cpp:228-238)
    T = q.SemiSparseSym3DTensor((naux, nao, nao), [[0]] + [[i] for i in range(1, nao)])
    with pytest.raises(IndexError):
        T[0, 3]  # unscreened pair (std::out_of_range, cpp:210)
    # NaN-filled until written (cpp:160)
    assert np.isnan(T.unique_dense_data).all()
    view = T.mut_unique_dense_data
    view[:, :] = 1.5
    assert (T.unique_dense_data == 1.5).all()
    with pytest.raises(RuntimeError):
        q.transform_integral(T, np.zeros((nao + 1, 2)), np.eye(nao), np.eye(naux), 1e-5)  # nao mismatch


def test_medium_properties(q):
    """Larger than the oracle can finish quickly: property checks instead of a reference value.
    (1) symmetry; (2) K3 against a torch FP64 matmul of the library's own X; (3) aux-slab additivity
    (the sharded path: sum of partial Gram matrices over P slabs equals the full one)."""
    import torch

    p = syn.make_chain_problem(n_units=14, ao_per_unit=24, n_end=5, naux=1100, frag_halfwidth=2, n_bath=40,
                               sigma=0.94, seed=105, n_frag=8)
    T = q.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable)
    T.whiten(p.L_PQ)
    TA = p.TAs[6]
    nmo = TA.shape[1]
    npair = nmo * (nmo + 1) // 2
    G = q.transform_integral(T, TA, p.S_abs, None, p.MO_coeff_epsilon)
    assert G.shape == (npair, npair)
    assert np.array_equal(G, G.T)
    # X of the same fragment, via the step-wise entry points
    A = q.get_AO_per_MO(TA, p.S_abs, p.MO_coeff_epsilon)
    X = q.contract_with_TA_2nd_to_sym_dense(q.contract_with_TA_1st(TA, T, A), TA)
    Xt = torch.from_numpy(np.ascontiguousarray(X)).cuda()
    Gt = (Xt.T @ Xt).cpu().numpy()
    assert np.abs(G - Gt).max() <= TOL
    # slabs
    cut = 512
    G1 = q.transform_integral(T, TA, p.S_abs, None, p.MO_coeff_epsilon, p_range=(0, cut))
    G2 = q.transform_integral(T, TA, p.S_abs, None, p.MO_coeff_epsilon, p_range=(cut, p.naux))
    assert np.abs(G1 + G2 - G).max() <= TOL


def test_pipelined_calls_with_pinned_outputs(q):
    """Back-to-back asynchronous transforms (double-buffered slots / G) give the same answers as
    synchronous ones."""
    p = syn.make_chain_problem(n_units=8, ao_per_unit=10, n_end=2, naux=160, frag_halfwidth=1, n_bath=10,
                               sigma=0.6, seed=106)
    T = q.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable)
    L = q.GPU_MatrixHandle(p.L_PQ)
    ref = [q.transform_integral(T, TA, p.S_abs, L, p.MO_coeff_epsilon).copy() for TA in p.TAs]
    outs = []
    for TA in p.TAs:
        nmo = TA.shape[1]
        npair = nmo * (nmo + 1) // 2
        o = _lib.pinned_empty((npair, npair))
        q.transform_integral(T, TA, p.S_abs, L, p.MO_coeff_epsilon, out=o, synchronize=False)
        outs.append(o)
    q.default_context().synchronize()
    for o, r in zip(outs, ref):
        assert np.array_equal(o, r)
    for o in outs:
        _lib.pinned_free(o)


def test_whitened_slab_staging_matches_oracle(q):
    """Aux-sharded staging on ONE device: every slab tensor holds only its rows of L^-1 (P|mu nu) (streamed in
    column chunks through scratch); the partial (ij|kl) of the slabs add up to the oracle's result."""
    from quemb_b200.sharded import aux_slabs

    p = syn.make_chain_problem(n_units=10, ao_per_unit=12, n_end=3, naux=333, frag_halfwidth=1, n_bath=20, sigma=0.7,
                               cond=1e8, seed=107)
    L = q.GPU_MatrixHandle(p.L_PQ)
    full = q.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable)
    full.whiten(L)
    W = full.unique_dense_data
    for world in (2, 3):
        slabs = aux_slabs(p.naux, world)
        Ts = [q.SemiSparseSym3DTensor.whitened_slab(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable, L,
                                                    s, chunk_cols=1000) for s in slabs]
        for (p0, p1), T in zip(slabs, Ts):
            assert T.whitened and T.shape == (p1 - p0, p.nao, p.nao)
            assert np.array_equal(T.unique_dense_data, W[p0:p1])  # same kernel, same rows: bit-identical
        TA = p.TAs[len(p.TAs) // 2]
        G = sum(q.transform_integral(T, TA, p.S_abs, None, p.MO_coeff_epsilon) for T in Ts)
        assert np.abs(G - _oracle_G(p, TA)).max() <= TOL
    with pytest.raises(RuntimeError):
        q.SemiSparseSym3DTensor.whitened_slab(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable, L, (8, 64))


def test_aux_sharded_build_two_gpus(q, tmp_path):
    """The NCCL reduce-sum path (north_star: oversize fragments shard P over the GPUs): torchrun, one rank per
    GPU, every fragment's (ij|kl) summed onto its owner rank and compared with the oracle there."""
    import os
    import subprocess
    import sys
    from pathlib import Path

    if _lib.device_count() < 2:
        pytest.skip("needs 2 GPUs (run under gpurun --gpus 2)")
    root = Path(__file__).resolve().parent.parent
    r = subprocess.run(
        [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
         "--master-port", "29641", str(root / "tests" / "mgpu_sharded_worker.py")],
        capture_output=True, text=True, timeout=600, env=dict(os.environ, PYTHONPATH=str(root)))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "SHARDED_OK" in r.stdout


def test_octane_full_size_vs_compiled_reference(q):
    """BASELINE.json configs[1] at full size (nao 202, naux 700, BE3-like fragments): the B200 path against the
    reference's own compiled extension (oracle/_ref, built by oracle/Makefile) -- screening pattern bit-exact,
    (ij|kl) <= 1e-10 -- for the smallest and the largest fragment, per-fragment solve and whitened tensor."""
    ref = oracle.reference_module()
    if ref is None:
        pytest.skip("oracle/_ref not built (needs /root/reference at build time)")
    p = syn.octane_like()
    T_ref = ref.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), [x.tolist() for x in p.exch_reachable])
    T = q.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable)
    Tw = q.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable)
    L = q.GPU_MatrixHandle(p.L_PQ)
    Tw.whiten(L)
    nmos = [t.shape[1] for t in p.TAs]
    for f in (int(np.argmin(nmos)), int(np.argmax(nmos))):
        TA = p.TAs[f]
        assert q.get_AO_per_MO(TA, p.S_abs, p.MO_coeff_epsilon) == [list(a) for a in ref.get_AO_per_MO(TA, p.S_abs, p.MO_coeff_epsilon)]
        G_ref = ref.transform_integral(T_ref, TA, p.S_abs, p.L_PQ, p.MO_coeff_epsilon)
        G = q.transform_integral(T, TA, p.S_abs, L, p.MO_coeff_epsilon)
        assert G.shape == G_ref.shape
        assert np.abs(G - G_ref).max() <= TOL
        Gw = q.transform_integral(Tw, TA, p.S_abs, None, p.MO_coeff_epsilon)
        assert np.abs(Gw - G_ref).max() <= TOL


@pytest.fixture(params=["stream", "rows"])
def jk_kernel(request, monkeypatch):
    """Both J/K kernels: the single-pass streaming kernel (jk_stream.cu, default for nmo <= 256) and the chunked
    row/column kernel (jk.cu, larger fragments); the library reads QEMB_JK_KERNEL on every call."""
    monkeypatch.setenv("QEMB_JK_KERNEL", request.param)
    return request.param


@pytest.mark.parametrize("name", ["chain_small", "lattice_small"])
def test_jk_consumer_on_resident_eri(q, name, jk_kernel):
    """Fused consumer: J/K and the HF-in-HF two-electron sum from the device-resident (ij|kl)
    against the CPU restatements applied to the REFERENCE's golden G."""
    from oracle import consumer_oracle as co
    from quemb_b200 import consumer

    g = Golden(name)
    T = _tensor(q, g)
    L = q.GPU_MatrixHandle(g.L)
    rng = np.random.default_rng(11)
    for f in range(g.n_frag):
        fr = g.frag(f)
        nmo = fr["TA"].shape[1]
        ws = q.transform_integral_resident(T, fr["TA"], g.S_abs, L, g.eps)
        npair = nmo * (nmo + 1) // 2
        assert np.abs(ws.G(npair) - fr["G"]).max() <= TOL
        C = np.linalg.qr(rng.standard_normal((nmo, nmo)))[0][:, : max(1, nmo // 3)]
        D = 2.0 * C @ C.T
        vj, vk = consumer.dot_eri_dm(ws, D)
        vj_ref, vk_ref = co.dot_eri_dm(fr["G"], D)
        assert np.abs(vj - vj_ref).max() <= 1e-10 and np.abs(vk - vk_ref).max() <= 1e-10
        vj2, vk2 = consumer.dot_eri_dm(ws, D)
        assert np.array_equal(vj, vj2) and np.array_equal(vk, vk2)  # deterministic (no atomics)
        n_frag = max(1, nmo // 2)
        e2 = consumer.ebe_hf_two_electron(ws, 0.5 * D, n_frag)
        assert np.abs(e2 - co.ebe_hf_e2(fr["G"], 0.5 * D, n_frag)).max() <= 1e-10
        # get_veff (helper.py:28-69) with an AO density / overlap / mean-field potential of matching shapes
        S = np.eye(g.nao) + 0.01 * g.S_abs
        dm = S @ fr["TA"] @ D @ fr["TA"].T @ S
        hf_veff = 0.5 * (g.S_abs + g.S_abs.T)
        Veff, Veff0 = consumer.get_veff(ws, dm, S, fr["TA"], hf_veff)
        ST = S @ fr["TA"]
        P_ = ST.T @ dm @ ST
        vj_r, vk_r = co.dot_eri_dm(fr["G"], 0.5 * (P_ + P_.T))
        assert np.abs(Veff - (fr["TA"].T @ hf_veff @ fr["TA"] - (vj_r - 0.5 * vk_r))).max() <= 1e-9 * max(1.0, np.abs(vj_r).max())
    with pytest.raises(RuntimeError):
        ws.jk(np.triu(np.ones((nmo, nmo))))  # not symmetric


def test_jk_multi_chunk_matches_torch(q, jk_kernel):
    """nmo large enough for several shared-memory chunks per column (npair > 12288): J/K against a torch FP64
    evaluation from the library's own resident G."""
    import torch

    p = syn.make_chain_problem(n_units=14, ao_per_unit=24, n_end=5, naux=300, frag_halfwidth=2, n_bath=60,
                               sigma=0.94, seed=109, n_frag=8)
    T = q.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable)
    T.whiten(p.L_PQ)
    TA = p.TAs[6]
    nmo = TA.shape[1]
    npair = nmo * (nmo + 1) // 2
    assert npair > 12288
    ws = q.transform_integral_resident(T, TA, p.S_abs, None, p.MO_coeff_epsilon)
    rng = np.random.default_rng(5)
    D = rng.standard_normal((nmo, nmo))
    D = D + D.T
    vj, vk = ws.jk(D)
    ms, nbytes = ws.jk_time()
    assert ms > 0 and nbytes == 8.0 * npair * npair
    Gt = torch.from_numpy(ws.G(npair)).cuda()
    hi, lo = np.tril_indices(nmo)
    idx = np.zeros((nmo, nmo), dtype=np.int64)
    idx[hi, lo] = np.arange(npair)
    idx[lo, hi] = np.arange(npair)
    it = torch.from_numpy(idx).cuda()
    Dt = torch.from_numpy(D).cuda()
    # J: packed weights; K: row-block evaluation to bound memory
    w = torch.where(torch.from_numpy(hi == lo).cuda(), 1.0, 2.0) * Dt[torch.from_numpy(hi).cuda(), torch.from_numpy(lo).cuda()]
    vj_ref = (Gt @ w)[it].cpu().numpy()
    assert np.abs(vj - vj_ref).max() <= 1e-10 * max(1.0, np.abs(vj_ref).max())
    vk_ref = np.zeros((nmo, nmo))
    for i in range(0, nmo, 16):  # K[i,k] = sum_jl G[idx[i,j], idx[k,l]] D[j,l]
        rows = it[i:i + 16].reshape(-1)                                   # (16*nmo)
        blk = Gt[rows][:, it.reshape(-1)].reshape(-1, nmo, nmo, nmo)      # [i, j, k, l]
        vk_ref[i:i + 16] = torch.einsum("ijkl,jl->ik", blk, Dt).cpu().numpy()
    assert np.abs(vk - vk_ref).max() <= 1e-10 * max(1.0, np.abs(vk_ref).max())


def test_jk_stream_kernel_all_block_structures(q, monkeypatch):
    """jk_stream.cu compiles the block structure of every 16-row stage in (QL = ceil(nmo / 32) variants, partial
    last stages, a lone last row): fragment sizes around every boundary, the streaming kernel against the chunked
    kernel (itself checked against torch / the oracle above), bitwise repeatable over many launches (a stage
    released before its reads returned would show up here), and the fallback above 256 orbitals."""
    p = syn.make_chain_problem(n_units=14, ao_per_unit=24, n_end=5, naux=96, frag_halfwidth=2, n_bath=60,
                               sigma=0.94, seed=311, n_frag=8)
    T = q.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable)
    T.whiten(p.L_PQ)
    rng = np.random.default_rng(17)
    for nmo in (1, 2, 15, 16, 17, 31, 32, 33, 48, 63, 65, 97, 128, 129, 161, 176, 177, 201, 224, 225, 255, 256, 260):
        TA = np.linalg.qr(rng.standard_normal((p.nao, nmo)))[0]
        ws = q.transform_integral_resident(T, TA, p.S_abs, None, 0.0)
        D = rng.standard_normal((nmo, nmo))
        D = D + D.T
        monkeypatch.setenv("QEMB_JK_KERNEL", "stream")
        vj, vk = ws.jk(D)
        reps = 40 if nmo in (33, 201, 256) else 2
        for _ in range(reps):
            vj2, vk2 = ws.jk(D)
            assert np.array_equal(vj, vj2) and np.array_equal(vk, vk2), nmo
        monkeypatch.setenv("QEMB_JK_KERNEL", "rows")
        vj_r, vk_r = ws.jk(D)
        sj, sk = max(1.0, np.abs(vj_r).max()), max(1.0, np.abs(vk_r).max())
        assert np.abs(vj - vj_r).max() <= 1e-12 * sj and np.abs(vk - vk_r).max() <= 1e-12 * sk, nmo
        if nmo > 256:
            assert np.array_equal(vj, vj_r)  # both calls ran the chunked kernel


def test_dense_int_direct_DF_on_the_same_kernels(q):
    """the unscreened limit against a numpy restatement of integral_direct_DF
    (eri_onthefly.py:127-145)."""
    import scipy.linalg as sla

    from quemb_b200 import dense_df

    rng = np.random.default_rng(21)
    naux, nao, nmo = 90, 23, 9
    pqL = rng.standard_normal((naux, nao, nao)) / np.sqrt(naux)
    pqL = pqL + pqL.transpose(0, 2, 1)
    low = syn.metric_cholesky(naux, 1e4, rng)
    TA = np.linalg.qr(rng.standard_normal((nao, nmo)))[0]
    got = {}
    dense_df.integral_direct_DF(pqL, low, [TA], lambda i, G: got.__setitem__(i, G.copy()))
    Lij = np.einsum("Lpq,pi,qj->Lij", pqL, TA, TA).reshape(naux, -1)
    bb = sla.solve_triangular(low, Lij, lower=True)
    eri_nosym = (bb.T @ bb).reshape(nmo, nmo, nmo, nmo)
    hi, lo = np.tril_indices(nmo)
    eri4 = eri_nosym[hi, lo][:, hi, lo]  # restore("4", ...): pair index T(max)+min
    assert got[0].shape == eri4.shape
    assert np.abs(got[0] - eri4).max() <= TOL


def test_pipeline_stress_repeated_runs(q):
    """Regression for the premature mbarrier stage release (round 1): K3 on the same X many times must equal a
    cuBLAS FP64 X^T X entry by entry every time (the race hit ~1 row of 1 warp in 1 of 10^4 tiles, i.e. most runs
    at this size), and K1/K2a must be bitwise reproducible."""
    import torch

    p = syn.octane_like()
    T = q.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable)
    L = q.GPU_MatrixHandle(p.L_PQ)
    TA = p.TAs[3]
    nmo = TA.shape[1]
    npair = nmo * (nmo + 1) // 2
    A = q.get_AO_per_MO(TA, p.S_abs, p.MO_coeff_epsilon)
    B0 = S0 = Gt = None
    for rep in range(12):
        B = q.contract_with_TA_1st(TA, T, A)
        Bd = B.dense_data
        Sd = q.contract_with_TA_2nd_to_sym_dense(B, TA)
        if B0 is None:
            B0, S0 = Bd, Sd
        assert np.array_equal(Bd, B0) and np.array_equal(Sd, S0)
        ws = q.transform_integral_resident(T, TA, p.S_abs, L, p.MO_coeff_epsilon)
        if Gt is None:
            X = np.linalg.solve(np.tril(p.L_PQ), S0)
            Xt = torch.from_numpy(np.ascontiguousarray(X)).cuda()
            Gt = (Xt.T @ Xt).cpu().numpy()
        assert np.abs(ws.G(npair) - Gt).max() <= TOL, f"repetition {rep}"


def test_borderline_screening_entries_are_reported(q):
    """entries within 1e-12 of the threshold are counted instead of passing silently."""
    nao, eps = 6, 1e-5
    S = np.eye(nao)
    TA = np.zeros((nao, 2), order="F")
    TA[0, 0] = 1.0          # far above
    TA[1, 0] = eps          # exactly on the threshold: kept (>=), borderline
    TA[2, 0] = eps * (1 - 1e-9)  # 1e-14 below: dropped, borderline
    TA[3, 0] = 0.5 * eps    # clearly dropped
    TA[4, 1] = -eps * (1 + 1e-9)  # 1e-14 above in magnitude: kept, borderline
    ws = q.default_workspace(q.default_context())
    A = q.get_AO_per_MO(TA, S, eps, ws=ws)
    assert A == [[0, 1], [4]]
    assert ws.borderline == 3
    assert A == [a.tolist() for a in oracle.get_AO_per_MO(TA, S, eps)]


@pytest.mark.parametrize("name", ["chain_small", "asym_sources"])
def test_block_staging_replaces_the_python_scatter_loop(q, name):
    """(mu nu|P) arrives as libcint-style blocks integrals[i, j, P] per (shell, shell range) and is
    scattered on the device; the tensor must equal, bit for bit, the one the reference's loop (py:459-488) fills."""
    g = Golden(name)
    T = q.SemiSparseSym3DTensor((g.naux, g.nao, g.nao), [x.tolist() for x in g.exch])
    rng = np.random.default_rng(2)
    shell = 3  # AOs per fake shell
    for p0 in range(0, g.nao, shell):
        n_i = min(shell, g.nao - p0)
        for q0 in range(0, p0 + n_i, 2 * shell):  # ranges of two shells, like get_blocks' contiguous runs
            n_j = min(2 * shell, g.nao - q0)
            blk = rng.standard_normal((n_i, n_j, g.naux))  # unscreened / upper-triangle entries: must be ignored
            for i in range(n_i):
                for j in range(n_j):
                    key = indexing.ravel_symmetric(p0 + i, q0 + j)
                    if q0 + j <= p0 + i and key in g.offsets:
                        blk[i, j, :] = g.M[:, g.offsets[key]]
            T.scatter_block(blk, p0, q0)
    np.testing.assert_array_equal(T.unique_dense_data, g.M)
    fr = g.frag(0)
    assert np.abs(q.transform_integral(T, fr["TA"], g.S_abs, g.L, g.eps) - fr["G"]).max() <= TOL


def test_fused_peer_accumulate_two_gpus(q):
    """K3 epilogue adds its tiles into the owner rank's matrix over NVLink peer memory (no separate collective):
    torchrun, one rank per GPU, compared with the oracle on the owner of every fragment."""
    import os
    import subprocess
    import sys
    from pathlib import Path

    if _lib.device_count() < 2:
        pytest.skip("needs 2 GPUs (run under gpurun --gpus 2)")
    root = Path(__file__).resolve().parent.parent
    r = subprocess.run(
        [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
         "--master-port", "29642", str(root / "tests" / "mgpu_fused_worker.py")],
        capture_output=True, text=True, timeout=600, env=dict(os.environ, PYTHONPATH=str(root)))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "FUSED_OK" in r.stdout


def test_accumulate_mode_and_device_mirror_single_gpu(q):
    """out_is_device = 2 on one device: two slabs accumulated into one zeroed matrix + device mirror = the oracle's
    (ij|kl); exercises the red.add epilogue, qemb_device_zero / _mirror_lower / _to_host and the IPC allocator."""
    G, G_oracle = _accumulate_two_slabs(q)
    assert np.array_equal(G, G.T)
    assert np.abs(G - G_oracle).max() <= TOL


def _accumulate_two_slabs(q):
    import ctypes

    from quemb_b200._lib import c_dbl, c_i64, c_int, c_p, check, lib, ptr
    from quemb_b200.sharded import aux_slabs

    p = syn.make_chain_problem(n_units=10, ao_per_unit=12, n_end=3, naux=333, frag_halfwidth=1, n_bath=20, sigma=0.7,
                               cond=1e8, seed=111)
    ctx = q.default_context()
    ws = q.default_workspace(ctx)
    L = q.GPU_MatrixHandle(p.L_PQ)
    Ts = [q.SemiSparseSym3DTensor.whitened_slab(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable, L, s)
          for s in aux_slabs(p.naux, 2)]
    TA = p.TAs[5]
    nmo = TA.shape[1]
    npair = nmo * (nmo + 1) // 2
    dev, handle = c_p(), (ctypes.c_ubyte * 64)()
    check(lib().qemb_ipc_alloc(ctx._h, c_i64(8 * npair * npair), ctypes.byref(dev), handle))
    check(lib().qemb_device_zero(ctx._h, dev, c_i64(8 * npair * npair)))
    sc = q._screen_for(p.S_abs, ctx)
    for T in Ts:
        check(lib().qemb_fragment_transform(ws._h, T._h, None, sc._h, ptr(TA), c_i64(TA.shape[0]), c_i64(TA.shape[0]),
                                            c_i64(nmo), c_dbl(p.MO_coeff_epsilon), c_i64(0), c_i64(-1), dev, c_i64(npair),
                                            c_int(2), None))
    check(lib().qemb_device_mirror_lower(ctx._h, dev, c_i64(npair), c_i64(npair)))
    G = np.empty((npair, npair), order="F")
    check(lib().qemb_device_to_host(ctx._h, ptr(G), dev, c_i64(8 * npair * npair)))
    check(lib().qemb_ipc_free(ctx._h, dev))
    return G, _oracle_G(p, TA)


def test_hybrid_output_transfer_matches_full_square(q):
    """Host-bound (ij|kl): triangle + host mirror, the hybrid (tail columns whole over PCIe) and the full-square
    transfer must deliver bit-identical matrices."""
    p = syn.make_chain_problem(n_units=14, ao_per_unit=24, n_end=5, naux=200, frag_halfwidth=2, n_bath=40,
                               sigma=0.94, seed=112, n_frag=8)
    ctx = q.default_context()
    T = q.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable)
    T.whiten(p.L_PQ)
    TA = p.TAs[6]
    try:
        ctx.set_output_mode(False)
        G0 = q.transform_integral(T, TA, p.S_abs, None, p.MO_coeff_epsilon).copy()
        ctx.set_output_mode(True)
        outs = []
        for frac in (0.0, 0.07, 0.5, 1.0):
            ctx.set_full_column_fraction(frac)
            outs.append(q.transform_integral(T, TA, p.S_abs, None, p.MO_coeff_epsilon).copy())
    finally:
        ctx.set_output_mode(True)
        ctx.set_full_column_fraction(0.0)
    assert G0.shape[0] > 4096
    for G in outs:
        assert np.array_equal(G, G0)


# int8-slice K3 (opt-in): kept last in the file -- the rounded base-256 slicer, 5 / 6 slices, the NaN rule and the
# accumulating form were made after the round-1 GPU minutes were spent and were checked on the host only (restatement,
# host twin of the device slicer)
@pytest.mark.parametrize("slices,rel", [(8, 2e-13), (7, 2e-13), (6, 1e-11), (5, 1e-9)])
def test_k3_int8_slices_opt_in(q, slices, rel):
    """Opt-in K3 on the int8 tensor cores (k3_int8.cu): error-free split of X into 8-bit slices (balanced digits of the rounded column), exact int32 products,
    FP64 recombination.  Against the reference's golden (ij|kl) and against the FP64 K3 of this library, on the
    goldens and on a fragment large enough for many tiles and k chunks; the default (0 slices) is restored."""
    ctx = q.default_context(0)
    cases = []
    for name in ("chain_small", "lattice_small"):
        g = Golden(name)
        T = _tensor(q, g)
        L = q.GPU_MatrixHandle(g.L)
        for f in range(g.n_frag):
            fr = g.frag(f)
            cases.append((T, fr["TA"], g.S_abs, L, g.eps, fr["G"]))
    p = syn.make_chain_problem(n_units=10, ao_per_unit=16, n_end=3, naux=333, frag_halfwidth=2, n_bath=40,
                               sigma=0.94, seed=77, n_frag=6)
    Tp = q.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable)
    Lp = q.GPU_MatrixHandle(p.L_PQ)
    cases.append((Tp, p.TAs[3], p.S_abs, Lp, p.MO_coeff_epsilon, None))
    try:
        for T, TA, S_abs, L, eps, G_gold in cases:
            ctx.set_k3_int8_slices(0)
            G64 = q.transform_integral(T, TA, S_abs, L, eps)
            ctx.set_k3_int8_slices(slices)
            G8 = q.transform_integral(T, TA, S_abs, L, eps)
            scale = max(1.0, np.abs(G64).max())
            assert G8.shape == G64.shape
            assert np.abs(G8 - G64).max() <= rel * scale
            assert np.array_equal(G8, G8.T)
            if G_gold is not None:
                assert np.abs(G8 - G_gold).max() <= max(TOL, rel * scale)
            G8b = q.transform_integral(T, TA, S_abs, L, eps)
            assert np.array_equal(G8, G8b)  # integer products: bitwise repeatable
    finally:
        ctx.set_k3_int8_slices(0)


@pytest.mark.parametrize("slices", [5, 6, 7, 8])
def test_k3_int8_bit_exact_against_its_restatement(q, slices):
    """The int8-slice K3 is integer work plus a fixed-order FP64 recombination: given the X the device holds
    (qemb_fragment_get_sym after the transform), the numpy restatement oracle/k3_int8_oracle.py must reproduce the
    device result bit for bit."""
    import ctypes

    from oracle import k3_int8_oracle as ko
    from quemb_b200._lib import c_i64, lib, ptr

    ctx = q.default_context(0)
    ws = q.default_workspace(ctx)
    try:
        ctx.set_k3_int8_slices(slices)
        for name in ("chain_small", "lattice_small"):
            g = Golden(name)
            T = _tensor(q, g)
            L = q.GPU_MatrixHandle(g.L)
            for f in range(g.n_frag):
                fr = g.frag(f)
                G8 = q.transform_integral(T, fr["TA"], g.S_abs, L, g.eps, ws=ws)
                naux, npair = g.L.shape[0], G8.shape[0]
                X = np.empty((naux, npair), order="F")
                assert lib().qemb_fragment_get_sym(ws._h, ptr(X), c_i64(naux)) == 0
                assert np.array_equal(G8, ko.gram_int8(X, slices)), (name, f)
    finally:
        ctx.set_k3_int8_slices(0)


def test_k3_int8_poisoned_pairs_like_the_fp64_product(q):
    """A NaN in (mu nu|P) reaches some columns of X; the FP64 K3 then holds NaN in the rows and columns of those pairs.
    The int8 K3 must flag exactly the same entries (slice_kernel marks the pair, the epilogue writes NaN) and agree
    on all the others."""
    g = Golden("chain_small")
    ctx = q.default_context(0)
    M = np.array(g.M, order="F", copy=True)
    M[3, M.shape[1] // 2] = np.nan
    T = q.SemiSparseSym3DTensor(M, (g.naux, g.nao, g.nao), [x.tolist() for x in g.exch])
    L = q.GPU_MatrixHandle(g.L)
    fr = g.frag(0)
    try:
        ctx.set_k3_int8_slices(0)
        G64 = q.transform_integral(T, fr["TA"], g.S_abs, L, g.eps)
        ctx.set_k3_int8_slices(7)
        G8 = q.transform_integral(T, fr["TA"], g.S_abs, L, g.eps)
    finally:
        ctx.set_k3_int8_slices(0)
    assert np.isnan(G64).any()
    assert np.array_equal(np.isnan(G8), np.isnan(G64))
    ok = ~np.isnan(G64)
    assert np.abs(G8[ok] - G64[ok]).max(initial=0.0) <= 1e-12 * max(1.0, np.abs(G64[ok]).max(initial=0.0))


@pytest.mark.parametrize("slices,tol", [(7, 1e-13), (6, 1e-11)])
def test_k3_int8_accumulating_slabs(q, slices, tol):
    """out_is_device = 2 with the int8-slice K3: each row slab is sliced with its own exponents and its lower triangle
    added into the shared matrix (what the fused aux-sharded build does over NVLink); the sum stays at the accuracy
    of a single int8 product (host study: profiles/k3_int8_parity_study_r1.txt, slab columns)."""
    ctx = q.default_context()
    G64, G_oracle = _accumulate_two_slabs(q)
    try:
        ctx.set_k3_int8_slices(slices)
        G, _ = _accumulate_two_slabs(q)
    finally:
        ctx.set_k3_int8_slices(0)
    assert np.array_equal(G, G.T)
    assert np.abs(G - G64).max() <= tol * max(1.0, np.abs(G64).max())
    assert np.abs(G - G_oracle).max() <= TOL


def test_k3_int8_guaranteed_tolerance_picks_the_smallest_sufficient_slice_count(q):
    """Guaranteed-tolerance mode of the opt-in int8 K3: e_max of X is reduced on the device, the smallest S whose
    worst-case bound meets the caller's absolute tolerance runs (the FP64 kernel when none does), and the result is
    within that tolerance of the FP64 product -- on reference goldens."""
    g = Golden("chain_small")
    ctx = q.default_context(0)
    T = q.SemiSparseSym3DTensor(g.M, (g.naux, g.nao, g.nao), g.exch, ctx=ctx)
    L = q.GPU_MatrixHandle(g.L, ctx=ctx)
    ws = q.Workspace(ctx)
    fr = g.frag(0)
    G64 = q.transform_integral(T, fr["TA"], g.S_abs, L, g.eps, ws=ws)
    assert ws.k3_choice[0] == 0
    # X of this fragment as the device holds it (solved in place) -> the exponent the device must find
    from quemb_b200._lib import c_i64, check, lib, ptr

    nmo = fr["TA"].shape[1]
    X = np.empty((g.naux, nmo * (nmo + 1) // 2), order="F")
    check(lib().qemb_fragment_get_sym(ws._h, ptr(X), c_i64(g.naux)))
    m = np.abs(X).max(axis=0)
    emax = int(np.frexp(m[m > 0].max())[1])
    rows = -(-g.naux // 16) * 16
    try:
        seen = set()
        for tol in (1e-30, 1e-13, 1e-11, 1e-9, 1e-6):
            ctx.set_k3_int8_tolerance(tol)
            G = q.transform_integral(T, fr["TA"], g.S_abs, L, g.eps, ws=ws)
            S, e = ws.k3_choice
            want = next((s for s in (5, 6, 7, 8) if q.k3_int8_bound(rows, s) * 2.0 ** (2 * emax) <= tol), 0)
            assert e == emax and S == want, (tol, S, want, e, emax)
            seen.add(S)
            if S == 0:
                assert np.array_equal(G, G64)
            else:
                assert np.abs(G - G64).max() <= tol + 1e-14 * np.abs(G64).max(), (tol, S)
            assert np.abs(G - fr["G"]).max() <= max(1e-10, tol)
        assert 0 in seen and len(seen) >= 3  # the tolerance actually steers the choice
    finally:
        ctx.set_k3_int8_tolerance(0.0)
    G = q.transform_integral(T, fr["TA"], g.S_abs, L, g.eps, ws=ws)
    assert ws.k3_choice[0] == 0 and np.array_equal(G, G64)
