"""Host-side one-offs in front of the fragment path: AO-AO screening and the staging of (P|mu nu).

Restates the pure-index parts of the reference driver so that the product seam does not depend on QuEmb's
Python module for them, and replaces its Python scatter loop by device staging:

* :func:`get_AO_per_AO`            ``_get_AO_per_AO``            ``src/quemb/molbe/eri_sparse_DF.py:225-240``
* :func:`account_for_symmetry`     ``account_for_symmetry``      ``eri_sparse_DF.py:342-359``
* :func:`identify_contiguous_blocks` / :func:`get_blocks`        ``eri_sparse_DF.py:379-398, 1009-1037``
* :func:`conversions_AO_shell`     ``conversions_AO_shell``      ``eri_sparse_DF.py:243-267``
* :func:`get_sparse_P_mu_nu`       ``get_sparse_P_mu_nu``        ``eri_sparse_DF.py:410-494``

libcint stays on the host (north star): the integral blocks come from an injected ``aux_e2`` callable with the
signature of the reference's ``_aux_e2`` (``eri_sparse_DF.py:80-120``; default: that very function, which needs
PySCF).  Every ``(n_i, n_j, naux)`` block goes straight to the device through
``SemiSparseSym3DTensor.scatter_block`` -- one pinned-slot copy, one H2D transfer and one scatter kernel per
block, overlapped with the evaluation of the next block -- instead of one numpy column assignment per AO pair
(``eri_sparse_DF.py:476-488``).  No host copy of (P|mu nu) is ever made.
"""

from __future__ import annotations

from collections import defaultdict
from itertools import takewhile
from typing import Callable, Mapping, Sequence

import numpy as np

from . import eri_sparse_DF as q


def get_AO_per_AO(S_abs: np.ndarray, epsilon: float, TA: np.ndarray | None = None) -> dict[int, np.ndarray]:
    """``_get_AO_per_AO`` (``eri_sparse_DF.py:225-240``): ``R(mu) = {nu : S_abs[nu, mu] >= epsilon}`` (``>=``) for
    the source AOs ``mu``; all AOs without ``TA``, otherwise only those with
    ``max_i (S_abs |TA|)[mu, i] > epsilon`` (strict ``>``)."""
    S_abs = np.asarray(S_abs)
    if TA is None:
        sources = range(len(S_abs))
    else:
        sources = ((S_abs @ abs(np.asarray(TA))).max(axis=1) > epsilon).nonzero()[0]
    return {int(i_AO): (S_abs[:, i_AO] >= epsilon).nonzero()[0] for i_AO in sources}


def exch_lists(exch_reachable: Mapping[int, Sequence[int]], nao: int) -> list[np.ndarray]:
    """Dense per-AO lists the tensor constructor takes; a missing key is an empty list
    (``exch_reachable.get(mu, [])``, ``eri_sparse_DF.py:433-436``)."""
    empty = np.zeros(0, dtype=np.int64)
    return [np.asarray(exch_reachable.get(mu, empty), dtype=np.int64) for mu in range(nao)]


def account_for_symmetry(reachable: Mapping[int, Sequence[int]]) -> dict[int, list[int]]:
    """``eri_sparse_DF.py:342-359``: keep the ``q <= p`` prefix of every sorted list."""
    return {p: list(takewhile(lambda x, p=p: p >= x, sorted(int(v) for v in qs))) for p, qs in reachable.items()}


def identify_contiguous_blocks(X: Sequence[int]) -> list[tuple[int, int]]:
    """``eri_sparse_DF.py:1009-1037``: (start, one-past-end) index pairs of runs of consecutive integers."""
    if len(X) == 0:
        return []
    result, start = [], 0
    for i in range(1, len(X)):
        if X[i] - X[i - 1] > 1:
            result.append((start, i))
            start = i
    result.append((start, len(X)))
    return result


def get_blocks(reachable: Sequence[int]) -> list[tuple[int, int]]:
    """``eri_sparse_DF.py:379-398``: (first value, last value) of every contiguous run."""
    return [(reachable[a], reachable[b - 1]) for a, b in identify_contiguous_blocks(reachable)]


def conversions_AO_shell(mol) -> tuple[dict[int, list[int]], dict[int, int]]:
    """``eri_sparse_DF.py:243-267`` (``mol.ao_loc_nr()`` / ``mol.nbas``)."""
    offsets = mol.ao_loc_nr()
    shell_id_to_AO = {s: list(range(int(offsets[s]), int(offsets[s + 1]))) for s in range(mol.nbas)}
    AO_to_shell_id = {ao: s for s, aos in shell_id_to_AO.items() for ao in aos}
    return shell_id_to_AO, AO_to_shell_id


def _default_aux_e2() -> Callable:
    from quemb.molbe.eri_sparse_DF import _aux_e2  # noqa: PLC0415  (needs PySCF / libcint)

    return _aux_e2


def get_sparse_P_mu_nu(mol, auxmol, exch_reachable: Mapping[int, Sequence[int]], *, aux_e2: Callable | None = None,
                       ctx: "q.Context | Sequence[q.Context] | None" = None, check_nan: bool = True):
    """Screened (P|mu nu) staged directly in HBM; same loop nest as ``eri_sparse_DF.py:459-488`` down to the libcint
    call (per shell, per contiguous block of reachable shells, all auxiliary shells), the per-pair Python scatter
    replaced by one :meth:`SemiSparseSym3DTensor.scatter_block` per block.  ``ctx`` may be a sequence of contexts
    (one per device): every block is evaluated ONCE on the host and scattered to each device; a list of tensors is
    returned."""
    aux_e2 = aux_e2 or _default_aux_e2()
    nao, naux = int(mol.nao), int(auxmol.nao)
    many = isinstance(ctx, (list, tuple))
    ctxs = list(ctx) if many else [ctx]
    lists = exch_lists(exch_reachable, nao)
    results = [q.SemiSparseSym3DTensor((naux, nao, nao), lists, ctx=c) for c in ctxs]
    exch_reachable_unique = account_for_symmetry(exch_reachable)
    shell_id_to_AO, AO_to_shell_id = conversions_AO_shell(mol)

    shell_reachable_by_shell: dict[int, set[int]] = defaultdict(set)
    for k, v in exch_reachable_unique.items():
        shell_reachable_by_shell[AO_to_shell_id[int(k)]] |= {AO_to_shell_id[int(orb)] for orb in v}

    for i_shell, reach in shell_reachable_by_shell.items():
        for start_block, stop_block in get_blocks(sorted(reach)):
            integrals = np.asarray(
                aux_e2(mol, auxmol, intor="int3c2e",
                       shls_slice=(i_shell, i_shell + 1, start_block, stop_block + 1, 0, auxmol.nbas)),
                order="C")
            # rows: the AOs of i_shell; columns: the AOs of shells start_block .. stop_block; the device scatter keeps
            # nu <= mu and screened pairs only ("still ensure p <= q" / key_to_offset test of the reference)
            for result in results:
                result.scatter_block(integrals, shell_id_to_AO[i_shell][0], shell_id_to_AO[start_block][0])
    if check_nan:
        for result in results:
            n = result.count_nan()
            if n:
                raise AssertionError(f"{n} NaN entries left in (P|mu nu) after staging (eri_sparse_DF.py:492)")
    return results if many else results[0]
