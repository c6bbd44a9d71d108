"""Host side of ``SemiSparse3DTensor`` (``pyi:38-75``): the reference's three public constructors, the index structures
they build (``cpp:405-423``: running offsets, map key ``ravel_Fortran(i_MO, mu, nao)``), ``mut_dense_data``, ``__getitem__``
and ``__repr__`` -- against the reference's own compiled class where ``oracle/_ref`` is built, and against the rules
written out here otherwise.  No device is touched."""
from __future__ import annotations

import numpy as np
import pytest

from oracle import oracle
from quemb_b200 import eri_sparse_DF as q

SHAPE = (7, 5, 4)  # naux, nao, nmo
LISTS = [[0, 2, 3], [], [1, 4], [0, 1, 2, 4]]


def _data(n, seed=5):
    return np.asfortranarray(np.random.default_rng(seed).standard_normal((SHAPE[0], n)))


def test_constructor_from_lists_allocates_nan_and_builds_running_offsets():
    t = q.SemiSparse3DTensor(SHAPE, LISTS)
    n = sum(len(x) for x in LISTS)
    assert t.shape == SHAPE and t.size == 7 * 5 * 4 and t.nonzero_size == 7 * n
    assert t.dense_data.shape == (7, n) and np.isnan(t.dense_data).all()  # cpp:346
    assert t.AO_reachable_by_MO == LISTS
    flat = [(i, mu) for i, row in enumerate(LISTS) for mu in row]
    assert t.AO_reachable_by_MO_with_offsets == [
        [(flat.index((i, mu)), mu) for mu in row] for i, row in enumerate(LISTS)]  # cpp:467-482 order
    assert t.offsets == {i + mu * SHAPE[1]: k for k, (i, mu) in enumerate(flat)}  # key as in cpp:414
    assert "Semi-sparse 3D tensor, shape=(7, 5, 4)" in repr(t)


def test_mut_dense_data_is_a_live_view_and_dense_data_is_read_only():
    t = q.SemiSparse3DTensor(SHAPE, LISTS)
    t.mut_dense_data[:, 2] = 3.5
    assert (t.dense_data[:, 2] == 3.5).all() and np.isnan(t.dense_data[:, 1]).all()
    assert t.mut_dense_data is t.mut_dense_data
    with pytest.raises(ValueError):
        t.dense_data[0, 0] = 1.0
    # __getitem__ goes through the offsets map: a constructor-built tensor files (i_MO, mu) under i_MO + mu * nao, and
    # get_aux_vector(a, b) looks up a + b * nao (cpp:383-387) -- so the stored pair answers to (i_MO, mu)
    np.testing.assert_array_equal(t[0, 3], t.dense_data[:, 2])
    with pytest.raises(IndexError):
        t[1, 0]


def test_constructor_copies_the_matrix_and_accepts_the_expert_form():
    n = sum(len(x) for x in LISTS)
    D = _data(n)
    t = q.SemiSparse3DTensor(D, SHAPE, LISTS)
    D[0, 0] += 1.0
    assert t.dense_data[0, 0] == D[0, 0] - 1.0  # pybind11 copies numpy -> Eigen
    perm = np.random.default_rng(1).permutation(n)
    with_off, k = [], 0
    for row in LISTS:
        with_off.append([(int(perm[k + j]), mu) for j, mu in enumerate(row)])
        k += len(row)
    offsets = {mu + i * SHAPE[1]: off for i, row in enumerate(with_off) for off, mu in row}
    e = q.SemiSparse3DTensor(D, SHAPE, LISTS, with_off, offsets)
    assert e.AO_reachable_by_MO_with_offsets == with_off and e.offsets == offsets
    np.testing.assert_array_equal(e[4, 2], D[:, with_off[2][1][0]])
    with pytest.raises(TypeError):
        q.SemiSparse3DTensor(D, SHAPE)


def test_same_answers_as_the_reference_class():
    ref = oracle.reference_module()
    if ref is None:
        pytest.skip("oracle/_ref not built on this box")
    n = sum(len(x) for x in LISTS)
    D = _data(n)
    for args in ((SHAPE, LISTS), (D, SHAPE, LISTS)):
        a, b = q.SemiSparse3DTensor(*args), ref.SemiSparse3DTensor(*args)
        assert a.shape == tuple(b.shape) and a.size == b.size and a.nonzero_size == b.nonzero_size
        assert a.AO_reachable_by_MO == [list(x) for x in b.AO_reachable_by_MO]
        assert a.AO_reachable_by_MO_with_offsets == [[tuple(p) for p in row] for row in b.AO_reachable_by_MO_with_offsets]
        assert a.offsets == dict(b.offsets)
        np.testing.assert_array_equal(a.dense_data, b.dense_data)
        assert repr(a) == repr(b)
    a, b = q.SemiSparse3DTensor(D, SHAPE, LISTS), ref.SemiSparse3DTensor(D, SHAPE, LISTS)
    for i, row in enumerate(LISTS):
        for mu in row:
            np.testing.assert_array_equal(a[i, mu], b[i, mu])
