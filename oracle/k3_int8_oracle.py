"""CPU restatement of the opt-in int8-slice K3 (quemb_b200/csrc/k3_int8.cu) -- TEST INFRASTRUCTURE ONLY.

Only ``tests/`` may import this module; the product never does (there is no CPU fallback).

Reference semantics: K3 of the reference is the FP64 product G = X^T X (``eri_sparse_DF.cpp:620``).  The int8
variant computes the same matrix through an error-free split of X into 8-bit integer slices:

    r = X[:, j] 2^-e_j (|r| < 1),   N = rint(r 2^(8S-2)),   N = sum_{a<S} D_a 256^(S-1-a),  D_a in [-128, 127]
    X[:, j] = 2^e_j * ( sum_{a<S} 2^-(8a+6) D_a[:, j] + R[:, j] ),     |R| <= 2^-(8S-1)
    G_ij    = 2^(e_i+e_j) * sum_{d<S} 2^-(8d+12) * sum_{a+b=d} <D_a[:, i], D_b[:, j]>      (integer sums exact)

N is rounded to nearest (ties to even) and its digits are the BALANCED base-256 digits, taken from the low end
(``csrc/k3_int8_digits.h``): neither the remainder nor the digits below the first have a preferred sign, so the
dropped terms add up like a random walk.  (Round 1 measured a truncating base-128 slicer on the GPU; its digits all
carried the sign of x and its dropped products lined up with x_i x_j: 3e-13 with 7 slices where this scheme gives
3e-16, ``profiles/k3_int8_parity_study_r1.txt``.)

Every step below is the kernel's arithmetic in the kernel's order (integer products are exact, the FP64
recombination adds the slice orders smallest first, the final scaling is a power of two), so the GPU result has to
match this restatement BIT FOR BIT.  Parity with the reference itself is pinned through the FP64 K3 (goldens of the
reference's own extension): ``tests/test_gpu_parity.py::test_k3_int8_slices_opt_in``.
"""
from __future__ import annotations

import numpy as np


EXPO_POISON = -(1 << 30)


def slice_int8(X: np.ndarray, slices: int):
    """Exponent per column, the integer slices (float64 arrays holding integers in [-128, 127], most significant
    first) and the remainder R in units of 2^e_j (|R| <= 2^-(8 slices - 1)).  Columns holding a NaN or an Inf get zero
    digits (``gram_int8`` marks them)."""
    assert 1 <= slices <= 8, "N has to fit 63 bits"
    X = np.array(X, dtype=np.float64)
    X[:, ~np.isfinite(X).all(axis=0)] = 0.0
    m = np.abs(X).max(axis=0) if X.shape[0] else np.zeros(X.shape[1])
    _, e = np.frexp(m)                       # m = f 2^e, f in [0.5, 1); e = 0 for m = 0
    e = np.where(m > 0, e, 0).astype(np.int64)
    K = 8 * slices - 2
    Nf = np.rint(np.ldexp(X, (K - e)[None, :]))   # nearest, ties to even (the kernel: llrint)
    R = np.ldexp(X, -e[None, :]) - np.ldexp(Nf, -K)  # exact
    N = Nf.astype(np.int64)
    D = [None] * slices
    for a in range(slices - 1, -1, -1):
        d = ((N + 128) & 255) - 128          # N = d mod 256
        D[a] = d.astype(np.float64)
        N = (N - d) >> 8                     # exact
    assert not N.any()
    return e, D, R


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
        g = acc[d] * 2.0 ** -(8 * d + 12) + g
    G = np.ldexp(g, e[:, None] + e[None, :])
    bad = ~np.isfinite(np.asarray(X, dtype=np.float64)).all(axis=0)  # as the FP64 product: row and column poisoned
    G[bad, :] = np.nan
    G[:, bad] = np.nan
    return G


def truncation_bound(naux: int, slices: int) -> float:
    """Worst-case |G_int8 - X^T X| / 2^(e_i+e_j): remainders of both operands plus the dropped slice orders."""
    rem = 2.0 ** -(8 * slices - 1)
    dropped = sum((min(d, slices - 1) - max(0, d - slices + 1) + 1) * 128.0 ** 2 * 2.0 ** -(8 * d + 12)
                  for d in range(slices, 2 * slices - 1))
    return naux * (2.0 * rem * (1.0 + rem) + rem * rem + dropped)
