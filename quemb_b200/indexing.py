"""Index maps of the sparse-DF path (host side, exact integer arithmetic).

Mirrors the reference's index helpers so that every map the device consumes is
bit-identical to what the reference builds:

* ``gauss_sum`` / ``ravel_symmetric`` / ``n_symmetric`` / ``ravel_Fortran``:
  reference ``src/quemb/molbe/_cpp/indexers.hpp:68-86,130-133`` (numba twins in
  ``src/quemb/shared/helper.py:249-357``).
* ``unravel_symmetric``: the reference uses a double ``sqrt``
  (``indexers.hpp:89-96``); here it is exact integer arithmetic (``isqrt``),
  which agrees with the reference wherever the latter is exact.
* ``extract_unique``: ``indexers.hpp:149-162`` (prefix with nu <= mu, breaks at
  the first nu > mu, i.e. assumes sorted lists).
* ``unique_offsets`` / ``reachable_columns``: the column assignment of
  ``SemiSparseSym3DTensor`` (``eri_sparse_DF.cpp:248-279``).
* ``mo_offsets``: ``get_AO_reachable_by_MO_with_offset``
  (``eri_sparse_DF.cpp:467-482``).

Lists of lists are carried as CSR pairs ``(ptr, idx)`` of ``int64`` arrays; this
is the format that crosses the C ABI (``include/qemb_b200.h``).
"""

from __future__ import annotations

from math import isqrt
from typing import Iterable, Sequence

import numpy as np


def gauss_sum(n: int) -> int:
    return (n * (n + 1)) // 2


def ravel_symmetric(a: int, b: int) -> int:
    return gauss_sum(a) + b if a > b else gauss_sum(b) + a


def n_symmetric(n: int) -> int:
    return ravel_symmetric(n - 1, n - 1) + 1


def unravel_symmetric(i: int) -> tuple[int, int]:
    """Inverse of :func:`ravel_symmetric`, returned as ``(min, max)`` exactly
    like the C++ reference (``indexers.hpp:95``). Note that the numba twin in
    ``shared/helper.py:286`` returns ``(max, min)``; the C++ order is the one the
    hot path (K2a) relies on."""
    a = (isqrt(8 * i + 1) - 1) // 2
    b = i - gauss_sum(a)
    return (a, b) if a <= b else (b, a)


def ravel_eri_idx(a: int, b: int, c: int, d: int) -> int:
    return ravel_symmetric(ravel_symmetric(a, b), ravel_symmetric(c, d))


def unravel_eri_idx(i: int) -> tuple[int, int, int, int]:
    """Returns (a, b, c, d) with a >= b, c >= d, ab >= cd like the numba twin
    (``shared/helper.py:305-310``) that ``tests/test_indexing.py`` pins."""
    cd, ab = unravel_symmetric(i)
    b, a = unravel_symmetric(ab)
    d, c = unravel_symmetric(cd)
    return a, b, c, d


def ravel_Fortran(a: int, b: int, n_rows: int) -> int:
    return a + b * n_rows


# ---------------------------------------------------------------------------
# CSR helpers
# ---------------------------------------------------------------------------


def to_csr(lists: Sequence[Iterable[int]]) -> tuple[np.ndarray, np.ndarray]:
    """list[list[int]] -> (ptr[n+1], idx[nnz]) int64."""
    n = len(lists)
    ptr = np.zeros(n + 1, dtype=np.int64)
    arrs = [np.asarray(x, dtype=np.int64).reshape(-1) for x in lists]
    if n:
        ptr[1:] = np.cumsum([a.size for a in arrs])
    idx = np.concatenate(arrs) if n and ptr[-1] > 0 else np.zeros(0, dtype=np.int64)
    return ptr, np.ascontiguousarray(idx, dtype=np.int64)


def from_csr(ptr: np.ndarray, idx: np.ndarray) -> list[list[int]]:
    return [idx[ptr[i] : ptr[i + 1]].tolist() for i in range(len(ptr) - 1)]


def extract_unique(exch_reachable: Sequence[Sequence[int]]) -> list[list[int]]:
    """``indexers.hpp:149-162``."""
    out: list[list[int]] = []
    for mu, lst in enumerate(exch_reachable):
        row: list[int] = []
        for nu in lst:
            if nu > mu:
                break
            row.append(int(nu))
        out.append(row)
    return out


def unique_offsets(exch_reachable_unique: Sequence[Sequence[int]]) -> dict[int, int]:
    """key ``ravel_symmetric(mu, nu)`` -> column; counter runs over mu, then the
    unique list of mu (``eri_sparse_DF.cpp:248-258``)."""
    offsets: dict[int, int] = {}
    counter = 0
    for mu, lst in enumerate(exch_reachable_unique):
        for nu in lst:
            offsets[ravel_symmetric(mu, nu)] = counter
            counter += 1
    return offsets


def reachable_columns(
    exch_reachable: Sequence[Sequence[int]], offsets: dict[int, int]
) -> list[list[tuple[int, int]]]:
    """``exch_reachable_with_offsets`` (``eri_sparse_DF.cpp:260-279``). A pair
    whose key was never assigned maps to column 0, which is what the
    reference's ``unordered_map::operator[]`` default-inserts."""
    return [
        [(offsets.get(ravel_symmetric(mu, nu), 0), int(nu)) for nu in lst]
        for mu, lst in enumerate(exch_reachable)
    ]


def mo_offsets(AO_by_MO: Sequence[Sequence[int]]) -> list[list[tuple[int, int]]]:
    """``get_AO_reachable_by_MO_with_offset`` (``eri_sparse_DF.cpp:467-482``)."""
    out = []
    counter = 0
    for aos in AO_by_MO:
        row = []
        for ao in aos:
            row.append((counter, int(ao)))
            counter += 1
        out.append(row)
    return out
