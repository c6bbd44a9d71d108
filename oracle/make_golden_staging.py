"""oracle/make_golden_staging.py -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Generates ``tests/golden/staging_rules.npz``: outputs of the REFERENCE's own pure-Python index rules
(``_get_AO_per_AO`` ``eri_sparse_DF.py:225-240``, ``account_for_symmetry`` ``:342-359``, ``identify_contiguous_blocks``
``:1009-1037``, ``get_blocks`` ``:379-398``) on seeded inputs, including entries that sit exactly on the thresholds
(the ``>=`` / strict ``>`` pair).

The reference module itself cannot be imported here (it imports PySCF at the top), so the four function definitions
are lifted out of ``/root/reference/src/quemb/molbe/eri_sparse_DF.py`` with ``ast`` -- decorators and annotations
dropped, bodies untouched -- and executed against numpy.  Nothing of the reference is written into the repository
except the numbers it produced.  Run in the build container only:  python oracle/make_golden_staging.py
"""
from __future__ import annotations

import ast
import sys
from itertools import takewhile
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
REF = Path("/root/reference/src/quemb/molbe/eri_sparse_DF.py")
WANTED = ("_get_AO_per_AO", "account_for_symmetry", "identify_contiguous_blocks", "get_blocks",
          "conversions_AO_shell", "_traverse_reachable", "get_sparse_P_mu_nu")


def reference_functions(path: Path = REF, extra: dict | None = None) -> dict:
    tree = ast.parse(path.read_text())
    picked = []
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in WANTED:
            node.decorator_list = []
            node.returns = None
            for a in node.args.args + node.args.kwonlyargs:
                a.annotation = None
            picked.append(node)
    mod = ast.Module(body=picked, type_ignores=[])
    ast.fix_missing_locations(mod)
    import __future__
    import logging
    from collections import defaultdict
    from typing import Mapping, Sequence  # the bodies mention Sequence[AOIdx] inside cast(...)

    class Timer:  # stand-in for quemb.shared.helper.Timer (logging only)
        def __init__(self, *_a):
            pass

        def str_elapsed(self):
            return "n/a"

    ns = {"np": np, "cast": lambda _t, x: x, "Sequence": Sequence, "Mapping": Mapping, "AOIdx": int, "ShellIdx": int,
          "takewhile": takewhile, "defaultdict": defaultdict, "Timer": Timer, "logger": logging.getLogger("ref"),
          # hpp:75-79 / shared/helper.py ravel_symmetric
          "ravel_symmetric": lambda a, b: (a * (a + 1) // 2 + b) if a > b else (b * (b + 1) // 2 + a)}
    ns.update(extra or {})
    exec(compile(mod, str(path), "exec", flags=__future__.annotations.compiler_flag), ns)  # noqa: S102
    return ns


def staging_cases():
    """(name, S_abs, eps, TA or None) -- seeded, with exact-threshold entries."""
    from quemb_b200 import synthetic as syn

    out = []
    p = syn.make_chain_problem(n_units=6, ao_per_unit=5, n_end=1, naux=24, frag_halfwidth=1, n_bath=4, sigma=0.5, seed=11)
    out.append(("chain_all", p.S_abs, 1e-10, None))
    out.append(("chain_frag0", p.S_abs, 1e-3, p.TAs[0]))
    out.append(("chain_frag3", p.S_abs, 1e-6, p.TAs[3]))
    # exact thresholds: S entries equal to eps (kept by >=), a source row whose max equals eps (dropped by >)
    eps = 0.25
    S = np.array([[1.0, 0.25, 0.0, 0.0],
                  [0.25, 1.0, 0.25, 0.0],
                  [0.0, 0.25, 1.0, 0.125],
                  [0.0, 0.0, 0.125, 1.0]])
    TA = np.array([[1.0, 0.0], [0.0, 0.0], [0.0, 0.0], [0.0, -2.0]])
    # (S |TA|) = [[1, 0], [.25, 0], [0, .25], [0, 2]] -> row maxima 1, .25, .25, 2: rows 1 and 2 sit ON eps (dropped)
    out.append(("exact_thresholds", S, eps, TA))
    out.append(("exact_thresholds_all", S, eps, None))
    return out


class StubMol:
    """What get_sparse_P_mu_nu / conversions_AO_shell touch of a pyscf Mole: nao, nbas, ao_loc_nr()."""

    def __init__(self, shell_sizes):
        self.ao_loc = np.concatenate([[0], np.cumsum(shell_sizes)]).astype(np.int64)
        self.nbas = len(shell_sizes)
        self.nao = int(self.ao_loc[-1])

    def ao_loc_nr(self):
        return self.ao_loc


def staging_problem(seed: int, on_fly_frag: int | None):
    """A seeded molecule-shaped staging input: shells of 1/3/5 AOs, dense symmetric (mu nu|P) source with NaN on every
    unscreened pair (nothing unscreened may ever be copied), exch_reachable from the reference's own rule."""
    from quemb_b200 import synthetic as syn

    p = syn.make_chain_problem(n_units=5, ao_per_unit=9, n_end=0, naux=40, frag_halfwidth=1, n_bath=5, sigma=0.45,
                               seed=seed)
    shells = []
    while sum(shells) < p.nao:
        shells.append(min((1, 3, 5)[len(shells) % 3], p.nao - sum(shells)))
    rng = np.random.default_rng(seed + 1)
    dense = rng.standard_normal((p.nao, p.nao, p.naux))
    dense = dense + dense.transpose(1, 0, 2)
    return p, shells, dense


def run_reference_staging(fns, ref_ext, p, shells, dense, exch):
    mol, auxmol = StubMol(shells), StubMol([4] * (p.naux // 4))
    lists = [np.asarray(exch.get(mu, []), dtype=np.int64) for mu in range(p.nao)]
    screened = np.zeros((p.nao, p.nao), dtype=bool)
    for mu, lst in enumerate(lists):
        screened[mu, lst] = True
    src = np.where((screened | screened.T)[:, :, None], dense, np.nan)

    def aux_e2(mol_, auxmol_, intor="int3c2e", shls_slice=None):
        i0, i1, j0, j1, k0, k1 = shls_slice
        assert intor == "int3c2e" and (k0, k1) == (0, auxmol_.nbas)
        return src[mol_.ao_loc[i0]: mol_.ao_loc[i1], mol_.ao_loc[j0]: mol_.ao_loc[j1], :].copy()

    fns["_aux_e2"] = aux_e2
    fns["SemiSparseSym3DTensor"] = ref_ext.SemiSparseSym3DTensor
    T = fns["get_sparse_P_mu_nu"](mol, auxmol, exch)
    return np.asarray(T.unique_dense_data), src, lists


def to_csr(d: dict, n: int):
    ptr = np.zeros(n + 1, dtype=np.int64)
    idx = []
    present = np.zeros(n, dtype=np.int8)
    for mu in range(n):
        if mu in d:
            present[mu] = 1
            idx.extend(int(x) for x in d[mu])
        ptr[mu + 1] = len(idx)
    return ptr, np.asarray(idx, dtype=np.int64), present


def main() -> None:
    fns = reference_functions()
    out = {}
    names = []
    for name, S, eps, TA in staging_cases():
        names.append(name)
        reach = fns["_get_AO_per_AO"](S, eps, TA)
        reach = {int(k): np.asarray(v) for k, v in reach.items()}
        ptr, idx, present = to_csr(reach, len(S))
        uniq = fns["account_for_symmetry"](reach)
        uptr, uidx, _ = to_csr({int(k): v for k, v in uniq.items()}, len(S))
        out[f"{name}/S"] = S
        out[f"{name}/eps"] = eps
        out[f"{name}/TA"] = np.zeros((0, 0)) if TA is None else TA
        out[f"{name}/ptr"], out[f"{name}/idx"], out[f"{name}/present"] = ptr, idx, present
        out[f"{name}/uptr"], out[f"{name}/uidx"] = uptr, uidx
    seqs = [[1, 2, 3, 5, 6, 7, 9, 10], [4], [0, 1, 2, 3], [0, 2, 4, 6], [3, 4, 9, 10, 11, 20], []]
    for k, X in enumerate(seqs):
        out[f"blocks/{k}/X"] = np.asarray(X, dtype=np.int64)
        out[f"blocks/{k}/identify"] = np.asarray(fns["identify_contiguous_blocks"](X), dtype=np.int64).reshape(-1, 2)
        out[f"blocks/{k}/get_blocks"] = np.asarray(fns["get_blocks"](X), dtype=np.int64).reshape(-1, 2)
    # get_sparse_P_mu_nu itself (py:410-494), run with the reference's compiled SemiSparseSym3DTensor (oracle/_ref) and
    # a stub mol / _aux_e2: pins the block loop, the p <= q rule and the key_to_offset test of the staging
    from oracle import oracle

    ref_ext = oracle.reference_module()
    assert ref_ext is not None, "oracle/_ref not built"
    stage_names = []
    for name, seed, frag, eps in (("stage_all", 21, None, 1e-10), ("stage_onfly", 22, 2, 1e-4)):
        p, shells, dense = staging_problem(seed, frag)
        exch = fns["_get_AO_per_AO"](p.S_abs, eps, None if frag is None else p.TAs[frag])
        exch = {int(k): np.asarray(v) for k, v in exch.items()}
        M, src, lists = run_reference_staging(fns, ref_ext, p, shells, dense, exch)
        assert not np.isnan(M).any()
        ptr, idx, present = to_csr(exch, p.nao)
        stage_names.append(name)
        out[f"{name}/shells"] = np.asarray(shells, dtype=np.int64)
        out[f"{name}/src"] = src
        out[f"{name}/ptr"], out[f"{name}/idx"], out[f"{name}/present"] = ptr, idx, present
        out[f"{name}/M"] = M
        out[f"{name}/naux"] = p.naux
    out["stage_names"] = np.asarray(stage_names)
    out["names"] = np.asarray(names)
    out["n_block_cases"] = len(seqs)
    dst = ROOT / "tests" / "golden" / "staging_rules.npz"
    np.savez_compressed(dst, **out)
    print("wrote", dst, dst.stat().st_size, "bytes;", ", ".join(names + stage_names))


if __name__ == "__main__":
    main()
