"""CPU: the host-built schedules the persistent kernels walk (no device): the lower-triangular tile order of K3, the
k split of small K3 launches, and the dealing of gather-GEMM tiles to the persistent CTAs.  Parity of the kernels
themselves is the GPU suite's job; these are the invariants a wrong table would break silently (a tile computed twice,
or never; an empty k split; a tile dropped by the balancer)."""
from __future__ import annotations

import ctypes
import os
import subprocess
import sys

import numpy as np
import pytest

from quemb_b200 import _lib


def _gram_tiles(nt: int) -> np.ndarray:
    out = np.full((nt * (nt + 1) // 2 + 1, 2), -1, dtype=np.int32)
    n = _lib.lib().qemb_test_gram_tiles(nt, _lib.ptr(out))
    assert n == nt * (nt + 1) // 2
    return out[:n]


@pytest.mark.parametrize("nt", [0, 1, 2, 11, 12, 13, 25, 148, 199, 400])
def test_k3_tile_walk_visits_every_lower_tile_once(nt):
    t = _gram_tiles(nt)
    assert (t[:, 0] >= t[:, 1]).all() and (t >= 0).all() and (t < max(nt, 1)).all()
    assert len({(int(a), int(b)) for a, b in t}) == nt * (nt + 1) // 2


def test_k3_tile_walk_keeps_a_wave_inside_one_band_of_row_tiles():
    """Default band = 12 row tiles: 148 consecutive tiles (one wave of the persistent kernel) touch at most 12 + 1 row
    panels away from a band boundary and ~148 / 12 column panels."""
    t = _gram_tiles(199)
    rows_per_wave = [len(set(t[w: w + 148, 0].tolist())) for w in range(0, len(t) - 148, 148)]
    cols_per_wave = [len(set(t[w: w + 148, 1].tolist())) for w in range(0, len(t) - 148, 148)]
    assert np.median(rows_per_wave) <= 12 and max(rows_per_wave) <= 24
    assert np.median(cols_per_wave) <= 14


@pytest.mark.parametrize("group", [1, 7, 20, 28, 500])
def test_k3_tile_walk_with_other_band_sizes_in_a_fresh_process(group):
    """QEMB_K3_GROUP is read once per process (the traffic experiments of DESIGN 4.5 used 12 ... 48)."""
    code = ("import sys, numpy as np; sys.path.insert(0, %r); from quemb_b200 import _lib\n"
            "for nt in (1, 5, 37, 199):\n"
            "    out = np.full((nt * (nt + 1) // 2 + 1, 2), -1, dtype=np.int32)\n"
            "    n = _lib.lib().qemb_test_gram_tiles(nt, _lib.ptr(out))\n"
            "    t = out[:n]\n"
            "    assert n == nt * (nt + 1) // 2 and (t[:, 0] >= t[:, 1]).all() and t.min() >= 0 and t.max() < nt\n"
            "    assert len({(int(a), int(b)) for a, b in t}) == n\n"
            "print('ok')" % str(_lib.ROOT))
    r = subprocess.run([sys.executable, "-c", code], env=dict(os.environ, QEMB_K3_GROUP=str(group)), capture_output=True,
                       text=True)
    assert r.returncode == 0 and r.stdout.strip() == "ok", r.stderr


def test_k3_split_rule_never_leaves_an_empty_split():
    L = _lib.lib()
    ns, kc = ctypes.c_int(), ctypes.c_int()
    for sms in (1, 132, 148):
        for ntiles in list(range(0, 40)) + [100, 147, 148, 149, 1000]:
            for ktiles in list(range(0, 70)) + [170, 500, 513]:
                _lib.check(L.qemb_test_gram_splits(ntiles, ktiles, sms, ctypes.byref(ns), ctypes.byref(kc)))
                n, c = ns.value, kc.value
                assert n >= 1
                if ntiles <= 0 or ntiles >= sms or ktiles < 8:
                    assert n == 1  # enough tiles to fill the SMs, or too little k to split
                    continue
                assert (n - 1) * c < ktiles <= n * c, (ntiles, ktiles, n, c)  # every split owns at least one k tile
                assert c >= 4 or n == 1                                       # ... and at least 4 when it is split
                assert n * ntiles <= 2 * sms + 2 * ntiles                     # about two work items per SM, not more


@pytest.mark.parametrize("n,sms", [(0, 148), (1, 148), (147, 148), (148, 148), (149, 148), (1000, 148), (4097, 132), (50, 1)])
def test_tile_balancer_returns_a_permutation_and_levels_the_load(n, sms):
    rng = np.random.default_rng(n)
    cost = (rng.integers(1, 20, size=n) * rng.integers(1, 9, size=n)).astype(np.float32)
    order = np.full(max(n, 1), -1, dtype=np.int32)
    _lib.check(_lib.lib().qemb_test_balance_tiles(n, _lib.ptr(cost) if n else None, sms, _lib.ptr(order)))
    assert sorted(order[:n].tolist()) == list(range(n))
    if n > sms > 1:
        # CTA c runs tiles c, c + sms, ...: no CTA carries more than the mean plus one largest tile (the serpentine deal)
        loads = np.array([cost[order[c:n:sms]].sum() for c in range(sms)])
        assert loads.max() <= cost.sum() / sms + cost.max() + 1e-3
