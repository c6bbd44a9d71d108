"""TEST INFRASTRUCTURE ONLY (see oracle/__init__ rules in DESIGN.md section 5): CPU restatements of the reference's
consumers of the fragment ERIs, used to check quemb_b200.consumer / qemb_fragment_jk.

dot_eri_dm lives in a third-party dependency that is absent from /root/reference and from both boxes: PySCF
(pyscf>=2.0.0, pyproject.toml:20 of the reference; no lock file, no fixture of it in the reference's tests).  Its
published algorithm is restated twice, independently:
  * `dot_eri_dm`        -- the two contractions of PySCF's own 4-index branch (pyscf/scf/hf.py, dot_eri_dm:
                           vj = einsum('ijkl,xji->xkl', eri, dms); vk = einsum('ijkl,xjk->xil', eri, dms)) on the
                           tensor unpacked from the s4 matrix by fancy indexing;
  * `dot_eri_dm_loops`  -- explicit Python loops: `restore1_loops` rebuilds (ij|kl) element by element from the
                           s4 matrix the way ao2mo.restore(1, eri, n) defines it (pair index T(max)+min on both
                           sides), then the same two sums as four nested loops.  Small nmo only.
tests/test_oracle.py checks them against each other (nmo <= 12, symmetric and non-symmetric dm) and against J/K
identities that do not involve either unpacking (tr(D vj) = tr(D' vj') exchange symmetry, e2 of the reference's own
update_ebe_hf loop).  With no PySCF anywhere this is what "pinned" can mean for this row: two independent
restatements of the published formula + the reference's own line-by-line loop (`ebe_hf_e2`, pfrag.py:368-389)
that consumes the same matrices.  `get_veff` follows helper.py:28-69 line by line.
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


def restore1_loops(G: np.ndarray, nmo: int) -> np.ndarray:
    """ao2mo.restore(1, eri_s4, nmo) by definition: eri[i,j,k,l] = G[pair(i,j), pair(k,l)], explicit loops."""
    eri = np.zeros((nmo, nmo, nmo, nmo))
    for i in range(nmo):
        for j in range(nmo):
            ij = i * (i + 1) // 2 + j if i >= j else j * (j + 1) // 2 + i
            for k in range(nmo):
                for l in range(nmo):
                    kl = k * (k + 1) // 2 + l if k >= l else l * (l + 1) // 2 + k
                    eri[i, j, k, l] = G[ij, kl]
    return eri


def dot_eri_dm_loops(G: np.ndarray, dm: np.ndarray):
    """PySCF's 4-index branch of dot_eri_dm written out: vj[k,l] = sum_ij eri[i,j,k,l] dm[j,i],
    vk[i,l] = sum_jk eri[i,j,k,l] dm[j,k]."""
    nmo = dm.shape[0]
    eri = restore1_loops(G, nmo)
    vj, vk = np.zeros((nmo, nmo)), np.zeros((nmo, nmo))
    for i in range(nmo):
        for j in range(nmo):
            for k in range(nmo):
                for l in range(nmo):
                    vj[k, l] += eri[i, j, k, l] * dm[j, i]
                    vk[i, l] += eri[i, j, k, l] * dm[j, k]
    return vj, vk


def get_veff(G: np.ndarray, dm: np.ndarray, S: np.ndarray, TA: np.ndarray, hf_veff: np.ndarray):
    """src/quemb/molbe/helper.py:28-69."""
    ST = S @ TA
    P_ = np.linalg.multi_dot((ST.T, dm, ST))
    P_ = np.asarray(P_.real, dtype=np.float64)
    vj, vk = dot_eri_dm(np.asarray(G, dtype=np.float64), P_)
    Veff_ = vj - 0.5 * vk
    Veff0 = np.linalg.multi_dot((TA.T, hf_veff, TA))
    return Veff0 - Veff_, Veff0


ebe_hf_two_electron = ebe_hf_e2
