"""CPU parity study of the opt-in int8-slice K3 over the test matrix -- TEST INFRASTRUCTURE (uses oracle/).

The device kernel (quemb_b200/csrc/k3_int8.cu) is bit-exact against oracle/k3_int8_oracle.py given the X the device
holds (tests/test_gpu_parity.py::test_k3_int8_bit_exact_against_its_restatement), so the error of the int8 variant
against the FP64 product can be studied on the host: this script sweeps the metric condition number, the geometry
family and the row-slab (aux-sharded / accumulating) decomposition, and writes one JSON line per case.

    python tests/int8_parity_study.py > profiles/k3_int8_parity_study_r1.jsonl      (then --extra, --bench-workload: appended)

"Exact" is the 8-slice restatement (62 bits below every column's exponent: truncation below naux * 2^-61 of the
column scales, then FP64 rounding of the sum), cross-checked against long double on a sample of entries -- that
cross-check ("exact_vs_longdouble") is at the same time the error of the 8-slice variant itself.
"""
from __future__ import annotations

import json
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

from oracle import k3_int8_oracle as ko  # noqa: E402
from oracle import oracle as orc  # noqa: E402
from quemb_b200 import synthetic as sy  # noqa: E402


def fragment_X(p, f: int) -> np.ndarray:
    """X = L^-1 (P|ij) of fragment f, through the restatement of the reference path (cpp:739-751 up to the trsm)."""
    TA = p.TAs[f]
    A = orc.get_AO_per_MO(TA, p.S_abs, p.MO_coeff_epsilon)
    g = orc.contract_with_TA_1st(TA, p.unique_dense_data, p.exch_reachable, A)
    sym = orc.contract_with_TA_2nd_to_sym_dense(g, TA, A)
    return orc.trsm_lower(p.L_PQ, sym)


def longdouble_sample(X: np.ndarray, rng, n: int = 4000):
    npair = X.shape[1]
    i, j = rng.integers(0, npair, n), rng.integers(0, npair, n)
    Xl = X.astype(np.longdouble)
    return i, j, np.einsum("pk,pk->k", Xl[:, i], Xl[:, j])


def study_case(name: str, X: np.ndarray, rng, slabs=(1, 2, 4, 8), slices=(6, 7)) -> dict:
    naux, npair = X.shape
    G64 = X.T @ X
    Gx = ko.gram_int8(X, 8)
    i, j, ld = longdouble_sample(X, rng)
    e, _, _ = ko.slice_int8(X, 7)
    out = dict(case=name, naux=naux, npair=npair, max_abs_G=float(np.abs(Gx).max()),
               max_scale=float(np.ldexp(1.0, 2 * int(e.max()))),
               exact_vs_longdouble=float(np.abs(Gx[i, j] - ld).max()),
               fp64_vs_exact=float(np.abs(G64 - Gx).max()))
    for S in slices:
        out[f"bound_x_scale_{S}"] = ko.truncation_bound(naux, S) * out["max_scale"]
        for ns in slabs:
            # row slabs of X: the aux-sharded path (one slab per rank, summed on the owner) and the accumulating
            # path (qemb_fragment_transform with accumulate) would each slice their slab with its own exponents
            edges = np.linspace(0, naux, ns + 1).astype(int)
            G8 = np.zeros((npair, npair))
            for a, b in zip(edges[:-1], edges[1:]):
                G8 += ko.gram_int8(X[a:b], S)
            out[f"int8x{S}_slabs{ns}"] = float(np.abs(G8 - Gx).max())
    return out


def cases(quick: bool):
    conds = (1e2, 1e8, 1e10) if quick else (1e2, 1e6, 1e8, 1e10, 1e12)
    for cond in conds:
        if quick:
            yield f"chain cond {cond:.0e}", sy.make_chain_problem(
                n_units=4, ao_per_unit=6, n_end=2, naux=160, frag_halfwidth=1, n_bath=8, cond=cond), (0, 2)
        else:
            yield f"chain cond {cond:.0e}", sy.make_chain_problem(
                n_units=8, ao_per_unit=12, n_end=3, naux=400, frag_halfwidth=2, n_bath=20, cond=cond), (0, 3)
            yield f"lattice cond {cond:.0e}", sy.make_lattice_problem(
                n_side=3, ao_per_site=6, naux=360, n_frag=2, n_bath=16, cond=cond), (0, 1)


def run(quick: bool = False, out=sys.stdout) -> list[dict]:
    rng = np.random.default_rng(20261104)
    rows = []
    for name, p, frags in cases(quick):
        for f in frags:
            row = study_case(f"{name} frag {f} (nmo {p.TAs[f].shape[1]})", fragment_X(p, f), rng,
                             slabs=(1, 4) if quick else (1, 2, 4, 8))
            rows.append(row)
            if out is not None:
                print(json.dumps(row), file=out, flush=True)
    return rows


def adversarial(rng, naux: int = 400, npair: int = 1200):
    """Operands the generators above do not produce: magnitudes log-uniform over 8 decades INSIDE every column (the
    per-column exponent is set by one entry, most entries keep few significant digits), one dominant auxiliary row,
    and columns of one sign (every product x_i x_j positive: nothing cancels)."""
    X = rng.standard_normal((naux, npair)) * 10.0 ** (-8.0 * rng.random((naux, npair)))
    yield "heavy tail inside columns (8 decades)", X
    X = rng.standard_normal((naux, npair)) * 1e-4
    X[naux // 3] = rng.standard_normal(npair)
    yield "one dominant auxiliary row (1e4 : 1)", X
    yield "all entries positive, uniform (0, 1)", rng.random((naux, npair))


def run_extra(out=sys.stdout) -> list[dict]:
    """5 slices on the geometry families at one condition number, and 5 / 6 / 7 slices on the adversarial operands."""
    rng = np.random.default_rng(20261105)
    rows = []
    for name, p, frags in cases(False):
        if "1e+08" in name:
            rows.append(study_case(f"{name} frag {frags[1]}", fragment_X(p, frags[1]), rng, slabs=(1, 4), slices=(5,)))
    for name, X in adversarial(rng):
        rows.append(study_case(name, X, rng, slabs=(1, 4), slices=(5, 6, 7)))
    for row in rows:
        if out is not None:
            print(json.dumps(row), file=out, flush=True)
    return rows


def run_bench_workload(out=sys.stdout, frags=(31, 15), nsel: int = 1500) -> list[dict]:
    """Operands of the benchmark workload itself (C32H66-like, naux 2716): per fragment the `nsel` pairs with the
    largest columns of X plus `nsel` random ones (the error of an entry depends on its two columns only)."""
    p = sy.c32h66_like(n_frag=32)
    rng = np.random.default_rng(20261106)
    rows = []
    for f in frags:
        X = fragment_X(p, f)
        big = np.argsort(-np.abs(X).max(axis=0))[:nsel]
        rest = np.setdiff1d(np.arange(X.shape[1]), big)
        sel = np.sort(np.concatenate([big, rng.choice(rest, nsel, replace=False)]))
        name = (f"C32H66-like (bench workload) frag {f} (nmo {p.TAs[f].shape[1]}), naux {p.naux}, {2 * nsel} of {X.shape[1]} "
                f"pairs: the {nsel} largest columns + {nsel} random")
        rows.append(study_case(name, np.asfortranarray(X[:, sel]), rng, slabs=(1,), slices=(5, 6, 7)))
        if out is not None:
            print(json.dumps(rows[-1]), file=out, flush=True)
    return rows


if __name__ == "__main__":
    if "--extra" in sys.argv:
        run_extra()
    elif "--bench-workload" in sys.argv:
        run_bench_workload()
    else:
        run(quick="--quick" in sys.argv)
