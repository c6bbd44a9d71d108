"""CPU restatement of the opt-in int8-slice K3 (quemb_b200/csrc/k3_int8.cu) -- TEST INFRASTRUCTURE ONLY.

Only ``tests/`` may import this module; the product never does (there is no CPU fallback).

Reference semantics: K3 of the reference is the FP64 product G = X^T X (``eri_sparse_DF.cpp:620``).  The int8
variant computes the same matrix through an error-free split of X into 7-bit integer slices:

    X[:, j] = 2^e_j * sum_{a<S} 2^-(7a+6) D_a[:, j] + R[:, j],     |R| <= 2^(e_j - 7S),   D_a int8 in [-64, 64]
    G_ij    = 2^(e_i+e_j) * sum_{d<S} 2^-(7d+12) * sum_{a+b=d} <D_a[:, i], D_b[:, j]>      (integer sums exact)

The digits are taken by ROUND TO NEAREST (ties to even, ``rint`` on both sides), not by truncation: the remainder
bound halves and, more important, the digits after the first have no preferred sign, so the dropped terms add up
like a random walk instead of all pointing the way of x_i x_j (measured: 35x smaller error at 7 slices,
``profiles/k3_int8_parity_study_r1.jsonl``).

Every step below is the kernel's arithmetic in the kernel's order (integer products are exact, the FP64
recombination adds the slice orders smallest first, the final scaling is a power of two), so the GPU result has to
match this restatement BIT FOR BIT.  Parity with the reference itself is pinned through the FP64 K3 (goldens of the
reference's own extension): ``tests/test_gpu_parity.py::test_k3_int8_slices_opt_in``.
"""
from __future__ import annotations

import numpy as np


def slice_int8(X: np.ndarray, slices: int):
    """Exponent per column, the integer slices (as float64 arrays holding integers in [-64, 64]) and the remainder
    in units of the last digit (|r| <= 1/2)."""
    X = np.array(X, dtype=np.float64)
    X[:, ~np.isfinite(X).all(axis=0)] = 0.0  # poisoned pairs (NaN / Inf): zero digits; gram_int8 marks them
    m = np.abs(X).max(axis=0) if X.shape[0] else np.zeros(X.shape[1])
    _, e = np.frexp(m)                       # m = f 2^e, f in [0.5, 1); e = 0 for m = 0
    e = np.where(m > 0, e, 0).astype(np.int64)
    r = np.ldexp(X, -e[None, :])             # exact, |r| < 1
    D, scale = [], 64.0
    for _ in range(slices):
        t = np.rint(r * scale)                # nearest, ties to even (the kernel: rint = cvt.rni.f64.f64)
        r = r * scale - t                    # exact
        D.append(t)
        scale = 128.0
    return e, D, r


def gram_int8(X: np.ndarray, slices: int) -> np.ndarray:
    """G = X^T X the way k3_int8.cu computes it."""
    e, D, _ = slice_int8(X, slices)
    npair = X.shape[1]
    acc = [np.zeros((npair, npair)) for _ in range(slices)]
    for a in range(slices):
        for b in range(slices - a):
            acc[a + b] += D[a].T @ D[b]      # integers below 2^31: exact in float64 whatever the summation order
    for d in range(slices):
        assert np.abs(acc[d]).max(initial=0.0) < 2.0 ** 31, "int32 accumulator of the kernel would overflow"
    g = np.zeros((npair, npair))
    for d in range(slices - 1, -1, -1):      # smallest terms first; product exact => same rounding as the kernel's fma
        g = acc[d] * 2.0 ** -(7 * d + 12) + g
    G = np.ldexp(g, e[:, None] + e[None, :])
    bad = ~np.isfinite(np.asarray(X, dtype=np.float64)).all(axis=0)  # as the FP64 product: row and column poisoned
    G[bad, :] = np.nan
    G[:, bad] = np.nan
    return G


def truncation_bound(naux: int, slices: int) -> float:
    """Worst-case |G_int8 - X^T X| / 2^(e_i+e_j): remainders of both operands plus the dropped slice orders."""
    rem = 2.0 ** -(7 * slices)
    dropped = sum((min(d, slices - 1) - max(0, d - slices + 1) + 1) * 64.0 ** 2 * 2.0 ** -(7 * d + 12)
                  for d in range(slices, 2 * slices - 1))
    return naux * (2.0 * rem * (1.0 + rem) + rem * rem + dropped)
