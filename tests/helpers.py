"""Shared helpers for the parity tests (loading the golden fixtures the reference produced)."""
from __future__ import annotations

from pathlib import Path

import numpy as np

GOLDEN = Path(__file__).resolve().parent / "golden"
CASES = ["chain_small", "lattice_small", "coarse_eps", "asym_sources"]


def csr_lists(ptr, idx):
    return [idx[ptr[i]: ptr[i + 1]].astype(np.int64) for i in range(len(ptr) - 1)]


class Golden:
    def __init__(self, name: str):
        self.name = name
        self.d = np.load(GOLDEN / f"{name}.npz")
        d = self.d
        self.nao, self.naux = int(d["nao"]), int(d["naux"])
        self.S_abs, self.L, self.M = d["S_abs"], d["L_PQ"], np.asfortranarray(d["M"])
        self.eps = float(d["eps"])
        self.exch = csr_lists(d["exch_ptr"], d["exch_idx"])
        self.uniq = csr_lists(d["uniq_ptr"], d["uniq_idx"])
        self.offsets = dict(zip(d["offset_keys"].tolist(), d["offset_vals"].tolist()))
        self.n_frag = int(d["n_frag"])

    def frag(self, f: int):
        d = self.d
        return dict(
            TA=np.asfortranarray(d[f"TA{f}"]),
            A=csr_lists(d[f"A_ptr{f}"], d[f"A_idx{f}"]),
            B=d[f"B{f}"], sym=d[f"sym{f}"], G=d[f"G{f}"],
        )
