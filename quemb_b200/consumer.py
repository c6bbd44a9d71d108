"""Consumers of the fragment ERIs, evaluated where (ij|kl) already lives.

The reference reads every fragment's ``(npair, npair)`` matrix back from HDF5 in ``BE.initialize``
(``src/quemb/molbe/mbe.py:1146-1170``) only to contract it with a density: ``get_veff``
(``src/quemb/molbe/helper.py:28-69``, pyscf ``scf.hf.dot_eri_dm``) for the fragment Fock matrix and the
two-electron sum of ``update_ebe_hf`` (``src/quemb/molbe/pfrag.py:368-389``) for the HF-in-HF check.  With the
J/K kernel of the library (``qemb_fragment_jk``) these need ``2 nmo^2`` doubles from the device instead of
``npair^2``; the same primitive serves every J/K build of a fragment SCF while G stays in HBM.

Same names and argument meaning as the reference functions; ``eri_`` is replaced by the workspace that holds
the resident (ij|kl) (``transform_integral_resident``).
"""

from __future__ import annotations

import numpy as np

from .eri_sparse_DF import Workspace


def dot_eri_dm(ws: Workspace, dm: np.ndarray) -> tuple[np.ndarray, np.ndarray]:
    """``pyscf.scf.hf.dot_eri_dm(eri, dm, hermi=1, with_j=True, with_k=True)`` on the resident ERIs."""
    return ws.jk(dm)


def get_veff(ws: Workspace, dm: np.ndarray, S: np.ndarray, TA: np.ndarray, hf_veff: np.ndarray):
    """``helper.get_veff`` (``helper.py:28-69``): Schmidt-space density ``P = (S TA)^T dm (S TA)``,
    ``Veff_ = vj - vk/2`` from the fragment ERIs, returns ``(TA^T hf_veff TA - Veff_, TA^T hf_veff TA)``."""
    ST = S @ TA
    P_ = np.asarray((ST.T @ dm @ ST).real, dtype=np.float64)
    P_ = 0.5 * (P_ + P_.T)  # hermi=1: the reference hands a Hermitian matrix to dot_eri_dm
    vj, vk = ws.jk(P_)
    Veff_ = vj - 0.5 * vk
    Veff0 = TA.T @ hf_veff @ TA
    return Veff0 - Veff_, Veff0


def ebe_hf_two_electron(ws: Workspace, rdm_hf: np.ndarray, n_frag: int) -> np.ndarray:
    """The ``e2`` vector of ``update_ebe_hf`` (``pfrag.py:368-389``, restricted case):
    ``e2[i] = sum_j Gij[tril] . eri[ij]`` with ``Gij = 2 D_ij D - D_i (x) D_j`` symmetrised, which is
    ``sum_j D_ij (2 J_ij - K_ij)`` for J, K built from ``D = rdm_hf``; ``i < n_frag``."""
    D = np.asarray(rdm_hf, dtype=np.float64)
    vj, vk = ws.jk(0.5 * (D + D.T))
    return np.einsum("ij,ij->i", D[:n_frag], 2.0 * vj[:n_frag] - vk[:n_frag])


def initialize_fragments(builder, TAs, MO_coeff_epsilon: float, hf_dm: np.ndarray, S: np.ndarray,
                         hf_veff: np.ndarray, rdm_hfs, n_frags, indices=None) -> dict[int, dict]:
    """The ERI-consuming part of ``BE.initialize`` (``mbe.py:1146-1170``) for ALL fragments in one device-resident
    pass: per fragment the transform leaves (ij|kl) in HBM and two J/K contractions are enqueued right behind it
    (``FragmentERIBuilder.build_jk``) -- ``get_veff`` for ``cons_fock`` (``helper.py:28-69``) and the two-electron
    sum of ``update_ebe_hf`` (``pfrag.py:368-389``) for the densities ``rdm_hfs[i]`` -- so only ``nmo x nmo``
    matrices leave the device and no ``eri_file`` round trip is needed.  ``builder`` is a staged
    :class:`quemb_b200.driver.FragmentERIBuilder`.  Returns ``{i: {"veff", "veff0", "e2", "vj", "vk"}}``."""
    idx = list(range(len(TAs))) if indices is None else list(indices)

    def densities(i):  # called per fragment from inside the pipeline: overlaps the device work of that fragment
        ST = S @ TAs[i]
        P_ = np.asarray((ST.T @ hf_dm @ ST).real, dtype=np.float64)
        D = np.asarray(rdm_hfs[i], dtype=np.float64)
        return [0.5 * (P_ + P_.T), 0.5 * (D + D.T)]

    jk = builder.build_jk(TAs, MO_coeff_epsilon, densities, indices=idx)
    out = {}
    for i in idx:
        (vj, vk), (vj2, vk2) = jk[i]
        TA = TAs[i]
        veff0 = TA.T @ hf_veff @ TA
        D = np.asarray(rdm_hfs[i], dtype=np.float64)
        n = int(n_frags[i])
        out[i] = {"veff": veff0 - (vj - 0.5 * vk), "veff0": veff0, "vj": vj, "vk": vk,
                  "e2": np.einsum("ij,ij->i", D[:n], 2.0 * vj2[:n] - vk2[:n])}
    return out
