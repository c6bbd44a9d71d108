"""TEST INFRASTRUCTURE ONLY (see oracle/__init__ rules in DESIGN.md section 5): CPU restatements of the reference's
consumers of the fragment ERIs, used to check quemb_b200.consumer / qemb_fragment_jk.

PARITY UNPINNED for dot_eri_dm: the algorithm lives in PySCF (pyscf>=2.0.0, pyproject.toml:20 of the reference;
not installed here or on the GPU box, no fixture of it in the reference's tests).  It is restated from its
published definition (pyscf/scf/hf.py `dot_eri_dm`: vj_ij = sum_kl (ij|kl) dm_lk, vk_il = sum_jk (ij|kl) dm_jk)
on the 4-index tensor unpacked from the s4 matrix.  `ebe_hf_e2` follows the reference's own loop line by line.
"""
import numpy as np


def unpack_s4(G: np.ndarray, nmo: int) -> np.ndarray:
    """(ij|kl) 4-index tensor from the npair x npair s4 matrix, pair index T(max)+min (indexers.hpp:75-79)."""
    idx = np.zeros((nmo, nmo), dtype=np.int64)
    for i in range(nmo):
        for j in range(nmo):
            hi, lo = max(i, j), min(i, j)
            idx[i, j] = hi * (hi + 1) // 2 + lo
    return G[idx[:, :, None, None], idx[None, None, :, :]]


def dot_eri_dm(G: np.ndarray, dm: np.ndarray):
    nmo = dm.shape[0]
    eri = unpack_s4(G, nmo)
    vj = np.einsum("ijkl,lk->ij", eri, dm)
    vk = np.einsum("ijkl,jk->il", eri, dm)
    return vj, vk


def ebe_hf_e2(G: np.ndarray, rdm_hf: np.ndarray, n_frag: int) -> np.ndarray:
    """src/quemb/molbe/pfrag.py:368-389 (restricted: unrestricted_fac = 2)."""
    jmax = rdm_hf.shape[0]
    e2 = np.zeros(n_frag)
    tril = np.tril_indices(jmax)
    for i in range(n_frag):
        for j in range(jmax):
            ij = i * (i + 1) // 2 + j if i > j else j * (j + 1) // 2 + i
            Gij = (2.0 * rdm_hf[i, j] * rdm_hf - np.outer(rdm_hf[i], rdm_hf[j]))[:jmax, :jmax]
            Gij[np.diag_indices(jmax)] *= 0.5
            Gij += Gij.T
            e2[i] += 0.5 * 2.0 * Gij[tril] @ G[ij]
    return e2
