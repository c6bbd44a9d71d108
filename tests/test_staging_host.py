"""Host-side staging rules against fixtures the REFERENCE's own functions produced
(``oracle/make_golden_staging.py`` lifts ``_get_AO_per_AO``, ``account_for_symmetry``, ``get_blocks`` and
``get_sparse_P_mu_nu`` out of ``/root/reference`` with ``ast`` and runs them; ``tests/golden/staging_rules.npz``)."""
from __future__ import annotations

import threading
import time
from pathlib import Path

import numpy as np
import pytest

from helpers import GOLDEN, csr_lists

from quemb_b200 import staging
from quemb_b200.eri_file import AsyncDatasetWriter, open_eri_file

D = np.load(GOLDEN / "staging_rules.npz")
RULE_CASES = [str(x) for x in D["names"]]
STAGE_CASES = [str(x) for x in D["stage_names"]]


def _exch_from_golden(name: str) -> dict[int, np.ndarray]:
    lists = csr_lists(D[f"{name}/ptr"], D[f"{name}/idx"])
    return {mu: lists[mu] for mu in range(len(lists)) if D[f"{name}/present"][mu]}


@pytest.mark.parametrize("name", RULE_CASES)
def test_get_AO_per_AO_matches_the_reference_rule(name):
    """``eri_sparse_DF.py:225-240``: keys = sources (strict ``>`` on the row maximum of ``S_abs |TA|``), lists = ``>=``."""
    S, eps, TA = D[f"{name}/S"], float(D[f"{name}/eps"]), D[f"{name}/TA"]
    got = staging.get_AO_per_AO(S, eps, None if TA.size == 0 else TA)
    want = _exch_from_golden(name)
    assert sorted(got) == sorted(want)
    for mu in want:
        assert np.array_equal(np.asarray(got[mu]), want[mu]), mu
    # symmetry-reduced lists (account_for_symmetry, py:342-359)
    uniq = staging.account_for_symmetry(got)
    want_u = csr_lists(D[f"{name}/uptr"], D[f"{name}/uidx"])
    for mu in want:
        assert uniq[mu] == want_u[mu].tolist()
    # dense lists handed to the tensor: a missing key is an empty list (py:433-436), never a KeyError
    lists = staging.exch_lists(got, len(S))
    assert len(lists) == len(S) and all(len(lists[mu]) == 0 for mu in range(len(S)) if mu not in want)


def test_exact_threshold_entries():
    """The ``>=`` / strict ``>`` pair on values that sit exactly on epsilon."""
    got = staging.get_AO_per_AO(D["exact_thresholds/S"], 0.25, D["exact_thresholds/TA"])
    assert sorted(got) == [0, 3]                      # rows 1 and 2 have max == eps: dropped by the strict '>'
    assert got[0].tolist() == [0, 1]                  # S[1, 0] == eps: kept by '>='
    assert got[3].tolist() == [3]


def test_block_helpers_match_the_reference():
    for k in range(int(D["n_block_cases"])):
        X = D[f"blocks/{k}/X"].tolist()
        assert np.array_equal(np.asarray(staging.identify_contiguous_blocks(X), dtype=np.int64).reshape(-1, 2),
                              D[f"blocks/{k}/identify"])
        assert np.array_equal(np.asarray(staging.get_blocks(X), dtype=np.int64).reshape(-1, 2),
                              D[f"blocks/{k}/get_blocks"])
    assert staging.account_for_symmetry({0: [0, 1, 2], 1: [0, 1, 2], 2: [0, 1, 2]}) == {0: [0], 1: [0, 1], 2: [0, 1, 2]}


def test_live_reference_rules_if_present():
    """In the build container the same functions are lifted from /root/reference again and compared directly."""
    if not Path("/root/reference/src/quemb/molbe/eri_sparse_DF.py").exists():
        pytest.skip("reference sources not on this box")
    import sys

    sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
    from oracle.make_golden_staging import reference_functions, staging_cases

    fns = reference_functions()
    for name, S, eps, TA in staging_cases():
        want = fns["_get_AO_per_AO"](S, eps, TA)
        got = staging.get_AO_per_AO(S, eps, TA)
        assert sorted(int(k) for k in want) == sorted(got), name
        for mu in got:
            assert np.array_equal(np.asarray(want[mu]), got[mu]), (name, mu)


class StubMol:
    """What the staging touches of a pyscf ``Mole``: ``nao``, ``nbas``, ``ao_loc_nr()``."""

    def __init__(self, shell_sizes):
        self.ao_loc = np.concatenate([[0], np.cumsum(shell_sizes)]).astype(np.int64)
        self.nbas = len(shell_sizes)
        self.nao = int(self.ao_loc[-1])

    def ao_loc_nr(self):
        return self.ao_loc


def stub_aux_e2(src):
    def aux_e2(mol, auxmol, intor="int3c2e", shls_slice=None):
        i0, i1, j0, j1, k0, k1 = shls_slice
        assert intor == "int3c2e" and (k0, k1) == (0, auxmol.nbas)
        return src[mol.ao_loc[i0]: mol.ao_loc[i1], mol.ao_loc[j0]: mol.ao_loc[j1], :].copy()

    return aux_e2


class NumpyTensor:
    """numpy statement of the C-ABI contract of ``qemb_tensor_create`` / ``qemb_tensor_scatter_block`` (column
    order ``cpp:248-258``; block rule ``include/qemb_b200.h``), so that the block loop of
    ``staging.get_sparse_P_mu_nu`` can be pinned on a box without a GPU; the device path has its own GPU test."""

    def __init__(self, shape, exch, ctx=None):
        self.naux, self.nao, _ = shape
        self.col = {}
        for mu, lst in enumerate(exch):
            for nu in lst:
                if nu > mu:
                    break
                self.col[(mu, int(nu))] = len(self.col)
        self.M = np.full((self.naux, len(self.col)), np.nan, order="F")

    def scatter_block(self, integrals, p0, q0):
        n_i, n_j, _ = integrals.shape
        for i in range(n_i):
            for j in range(n_j):
                mu, nu = p0 + i, q0 + j
                if nu > mu:
                    break
                c = self.col.get((mu, nu))
                if c is not None:
                    self.M[:, c] = integrals[i, j]

    def count_nan(self):
        return int(np.isnan(self.M).sum())


@pytest.mark.parametrize("name", STAGE_CASES)
def test_staging_block_loop_matches_reference_get_sparse_P_mu_nu(name, monkeypatch):
    """``eri_sparse_DF.py:410-494`` (run by the reference itself into its own compiled tensor -> golden ``M``)."""
    shells = D[f"{name}/shells"].tolist()
    src, naux = D[f"{name}/src"], int(D[f"{name}/naux"])
    mol, auxmol = StubMol(shells), StubMol([4] * (naux // 4))
    exch = _exch_from_golden(name)
    monkeypatch.setattr(staging.q, "SemiSparseSym3DTensor", NumpyTensor)
    T = staging.get_sparse_P_mu_nu(mol, auxmol, exch, aux_e2=stub_aux_e2(src))
    assert T.M.shape == D[f"{name}/M"].shape
    assert np.array_equal(T.M, D[f"{name}/M"])


def test_async_dataset_writer_orders_and_releases(tmp_path):
    """Background writer: datasets land with the reference's names / shape / dtype
    (``eri_sparse_DF.py:631``), futures complete, errors surface, close() flushes."""
    rng = np.random.default_rng(1)
    mats = [rng.standard_normal((n, n)) for n in (6, 10, 15)]
    mats = [m + m.T for m in mats]
    with open_eri_file(tmp_path / "eri.h5", "w") as f:
        w = AsyncDatasetWriter(f, max_pending=2)
        futs = [w.submit(f"f{i}", m) for i, m in enumerate(mats)]
        w.close()
        assert all(fu.done() and fu.result() == f"f{i}" for i, fu in enumerate(futs))
        assert w.bytes_written == sum(m.nbytes for m in mats)
    with open_eri_file(tmp_path / "eri.h5", "r") as f:
        for i, m in enumerate(mats):
            a = np.array(f.get(f"f{i}"))
            assert a.dtype == np.float64 and a.shape == m.shape and np.array_equal(a, m)

    class Slow:
        def __init__(self):
            self.names, self.active, self.max_active = [], 0, 0
            self.lock = threading.Lock()

        def create_dataset(self, name, data):
            with self.lock:
                self.active += 1
                self.max_active = max(self.max_active, self.active)
            time.sleep(0.01)
            if name == "bad":
                with self.lock:
                    self.active -= 1
                raise OSError("disk full")
            self.names.append(name)
            with self.lock:
                self.active -= 1

    s = Slow()
    with AsyncDatasetWriter(s) as w:
        futs = [w.submit(n, mats[0]) for n in ("a", "bad", "b")]
    assert s.names == ["a", "b"] and s.max_active == 1  # one writer, submission order
    assert futs[0].result() == "a" and futs[2].result() == "b"
    with pytest.raises(OSError):
        futs[1].result()
