"""Seeded synthetic inputs for the sparse-DF fragment ERI transform.

There is no PySCF / libcint on the build or GPU boxes, so parity tests and the
benchmark feed the path with inputs that have the *structure* of the real ones
(SURVEY.md §8d): AO centres on an alkane-like chain (or a jittered lattice),
a Gaussian absolute-overlap matrix ``S_abs``, AO-pair screening by the rule of
``_get_AO_per_AO`` (reference ``src/quemb/molbe/eri_sparse_DF.py:225-240``),
fragment orbitals = unit vectors on a contiguous AO block + orthonormal bath
tails (cf. ``src/quemb/molbe/pfrag.py:146-180``), and a Coulomb-metric
Cholesky factor with a prescribed condition number.

(P|mu nu) is generated as ``L @ Y`` with ``Y`` of O(1/sqrt(naux)) decaying
entries, i.e. the fitted quantities ``L^-1 (P|mu nu)`` are O(1) as they are for
real molecules; the resulting (ij|kl) are O(1) and a 1e-10 max-abs tolerance
means what it means in the reference's tests.
"""

from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np
import scipy.linalg

SEED = 20261104


@dataclass
class Problem:
    """Everything ``transform_integral`` consumes, for all fragments."""

    nao: int
    naux: int
    S_abs: np.ndarray  # (nao, nao) float64
    exch_reachable: list[np.ndarray]  # per AO mu: sorted int64 nu list
    unique_dense_data: np.ndarray  # (naux, n_unique) float64, Fortran order
    L_PQ: np.ndarray  # (naux, naux) float64 lower triangular
    TAs: list[np.ndarray]  # per fragment (nao, nmo) float64
    MO_coeff_epsilon: float = 1e-5
    AO_coeff_epsilon: float = 1e-10
    name: str = "synthetic"
    meta: dict = field(default_factory=dict)

    @property
    def n_unique(self) -> int:
        return self.unique_dense_data.shape[1]


def get_AO_per_AO(S_abs: np.ndarray, epsilon: float) -> list[np.ndarray]:
    """R(mu) = {nu : S_abs[nu, mu] >= eps}, ascending
    (reference ``eri_sparse_DF.py:237-240``, the ``TA is None`` branch)."""
    return [
        np.nonzero(S_abs[:, mu] >= epsilon)[0].astype(np.int64)
        for mu in range(S_abs.shape[0])
    ]


def unique_pairs(exch_reachable: list[np.ndarray]) -> tuple[np.ndarray, np.ndarray]:
    """(mu, nu) of every unique column in column order (mu ascending, then the
    nu <= mu prefix of R(mu)); reference ``eri_sparse_DF.cpp:248-258``."""
    mus, nus = [], []
    for mu, lst in enumerate(exch_reachable):
        pre = lst[: np.searchsorted(lst, mu, side="right")]
        mus.append(np.full(pre.size, mu, dtype=np.int64))
        nus.append(pre)
    if not mus:
        return np.zeros(0, np.int64), np.zeros(0, np.int64)
    return np.concatenate(mus), np.concatenate(nus)


def chain_centres(n_units: int, ao_per_unit: int, n_end: int, rng) -> np.ndarray:
    """1-D alkane-like chain: ``n_units`` CH2-like groups of ``ao_per_unit`` AOs
    one length unit apart, plus ``n_end`` extra AOs on each terminus."""
    x = []
    x += [(-0.4, 0.0, 0.0)] * n_end
    for k in range(n_units):
        x += [(float(k), 0.0, 0.0)] * ao_per_unit
    x += [(n_units - 1 + 0.4, 0.0, 0.0)] * n_end
    r = np.asarray(x, dtype=np.float64)
    r[:, 1:] += 0.15 * rng.standard_normal((len(r), 2))
    return r


def lattice_centres(n_side: int, ao_per_site: int, rng) -> np.ndarray:
    """3-D jittered cubic lattice (water-cluster-like)."""
    g = np.stack(
        np.meshgrid(*(np.arange(n_side, dtype=np.float64),) * 3, indexing="ij"), -1
    ).reshape(-1, 3)
    g = g + 0.2 * rng.standard_normal(g.shape)
    return np.repeat(g, ao_per_site, axis=0)


def gaussian_S_abs(r: np.ndarray, sigma: float) -> np.ndarray:
    d2 = ((r[:, None, :] - r[None, :, :]) ** 2).sum(-1)
    S = np.exp(-d2 / (2.0 * sigma * sigma))
    np.fill_diagonal(S, 1.0)
    return S


def metric_cholesky(naux: int, cond: float, rng, device_linalg: bool = False) -> np.ndarray:
    """cholesky(Q diag(logspace(0,-log10 cond)) Q^T, lower=True) with a random
    orthogonal Q — an RI-metric-like ill-conditioned SPD matrix.  ``device_linalg``: QR / products / Cholesky with
    torch on the current CUDA device (synthetic-data generation of the big timing cases only; naux = 8000 costs a
    minute per rank on a few host threads)."""
    A = rng.standard_normal((naux, naux))
    ev = np.logspace(0.0, -np.log10(cond), naux)
    if device_linalg:
        import torch

        Ad = torch.from_numpy(A).cuda()
        Qd, _ = torch.linalg.qr(Ad)
        PQd = (Qd * torch.from_numpy(ev).cuda()) @ Qd.T
        PQd = 0.5 * (PQd + PQd.T)
        Ld = torch.linalg.cholesky(PQd)
        L = np.asfortranarray(Ld.cpu().numpy())
        del Ad, Qd, PQd, Ld
        torch.cuda.empty_cache()
        return L
    Q, _ = np.linalg.qr(A)
    PQ = (Q * ev) @ Q.T
    PQ = 0.5 * (PQ + PQ.T)
    return np.asfortranarray(scipy.linalg.cholesky(PQ, lower=True))


def fragment_TA(
    r: np.ndarray,
    frag_aos: np.ndarray,
    n_bath: int,
    lam: float,
    rng,
) -> np.ndarray:
    """TA (nao, n_f + n_bath): unit vectors on the fragment AOs, then bath
    orbitals: orthonormalised random tails living on the environment AOs with
    exp(-d/lam) decay in the distance d to the fragment."""
    nao = len(r)
    n_f = len(frag_aos)
    TA = np.zeros((nao, n_f + n_bath), dtype=np.float64)
    TA[frag_aos, np.arange(n_f)] = 1.0
    if n_bath:
        env = np.setdiff1d(np.arange(nao), frag_aos)
        n_bath_eff = min(n_bath, env.size)
        d = np.sqrt(
            ((r[env, None, :] - r[None, frag_aos, :]) ** 2).sum(-1)
        ).min(axis=1)
        tails = rng.standard_normal((env.size, n_bath_eff)) * np.exp(-d / lam)[:, None]
        Qb, _ = np.linalg.qr(tails)
        TA[env, n_f : n_f + n_bath_eff] = Qb
        TA = TA[:, : n_f + n_bath_eff]
    return np.asfortranarray(TA)


def make_P_mu_nu(
    r: np.ndarray,
    sigma: float,
    exch_reachable: list[np.ndarray],
    L_PQ: np.ndarray,
    rng,
    whitened_scale: bool = True,
    matmul=None,
) -> np.ndarray:
    """unique_dense_data (naux, n_unique) = L @ Y, Y ~ N(0,1) exp(-d^2/4 sigma^2)/sqrt(naux)."""
    naux = L_PQ.shape[0]
    mus, nus = unique_pairs(exch_reachable)
    d2 = ((r[mus] - r[nus]) ** 2).sum(-1)
    decay = np.exp(-d2 / (4.0 * sigma * sigma))
    Y = rng.standard_normal((naux, mus.size)) * decay[None, :]
    if whitened_scale:
        Y /= np.sqrt(naux)
        M = (matmul or np.matmul)(np.tril(L_PQ), Y)  # `matmul`: optional accelerator for the big cases
    else:
        M = Y
    return np.asfortranarray(M)


def make_chain_problem(
    n_units: int = 8,
    ao_per_unit: int = 6,
    n_end: int = 1,
    naux: int = 96,
    frag_halfwidth: int = 1,
    n_bath: int | None = None,
    sigma: float = 0.94,
    lam: float = 1.0,
    cond: float = 1e8,
    n_frag: int | None = None,
    seed: int = SEED,
    MO_coeff_epsilon: float = 1e-5,
    AO_coeff_epsilon: float = 1e-10,
    name: str = "chain",
    matmul=None,
    device_linalg: bool = False,
) -> Problem:
    """Alkane-like chain. One fragment per unit (centre unit +- halfwidth units,
    clipped at the ends — edge fragments are smaller, like BE-n on a chain)."""
    rng = np.random.default_rng(seed)
    r = chain_centres(n_units, ao_per_unit, n_end, rng)
    nao = len(r)
    S_abs = gaussian_S_abs(r, sigma)
    exch = get_AO_per_AO(S_abs, AO_coeff_epsilon)
    L = metric_cholesky(naux, cond, rng, device_linalg=device_linalg)
    M = make_P_mu_nu(r, sigma, exch, L, rng, matmul=matmul)
    unit_of_ao = np.concatenate(
        [
            np.zeros(n_end, np.int64),
            np.repeat(np.arange(n_units), ao_per_unit),
            np.full(n_end, n_units - 1, np.int64),
        ]
    )
    centres = list(range(n_units))
    if n_frag is not None:
        centres = centres[:n_frag]
    TAs = []
    for c in centres:
        lo, hi = max(0, c - frag_halfwidth), min(n_units - 1, c + frag_halfwidth)
        frag_aos = np.nonzero((unit_of_ao >= lo) & (unit_of_ao <= hi))[0]
        nb = len(frag_aos) if n_bath is None else min(n_bath, len(frag_aos))
        TAs.append(fragment_TA(r, frag_aos, nb, lam, rng))
    return Problem(
        nao=nao,
        naux=naux,
        S_abs=np.asfortranarray(S_abs),
        exch_reachable=exch,
        unique_dense_data=M,
        L_PQ=L,
        TAs=TAs,
        MO_coeff_epsilon=MO_coeff_epsilon,
        AO_coeff_epsilon=AO_coeff_epsilon,
        name=name,
        meta=dict(
            n_units=n_units, ao_per_unit=ao_per_unit, n_end=n_end, sigma=sigma,
            lam=lam, cond=cond, seed=seed, frag_halfwidth=frag_halfwidth,
        ),
    )


def make_lattice_problem(
    n_side: int = 3,
    ao_per_site: int = 4,
    naux: int = 80,
    n_frag: int = 3,
    n_bath: int | None = None,
    sigma: float = 0.6,
    lam: float = 0.8,
    cond: float = 1e6,
    seed: int = SEED + 1,
    MO_coeff_epsilon: float = 1e-5,
    AO_coeff_epsilon: float = 1e-10,
    name: str = "lattice",
    matmul=None,
    n_sites: int | None = None,
) -> Problem:
    """Water-cluster-like 3-D lattice; a fragment = one site + its nearest sites."""
    rng = np.random.default_rng(seed)
    r = lattice_centres(n_side, ao_per_site, rng)
    nao = len(r)
    S_abs = gaussian_S_abs(r, sigma)
    exch = get_AO_per_AO(S_abs, AO_coeff_epsilon)
    L = metric_cholesky(naux, cond, rng)
    M = make_P_mu_nu(r, sigma, exch, L, rng, matmul=matmul)
    site_of_ao = np.repeat(np.arange(n_side**3), ao_per_site)
    site_r = r[::ao_per_site]
    TAs = []
    for c in range(n_frag):
        d = np.sqrt(((site_r - site_r[c]) ** 2).sum(-1))
        sites = np.nonzero(d < 1.3)[0] if n_sites is None else np.sort(np.argsort(d, kind="stable")[:n_sites])
        frag_aos = np.nonzero(np.isin(site_of_ao, sites))[0]
        nb = len(frag_aos) if n_bath is None else min(n_bath, len(frag_aos))
        TAs.append(fragment_TA(r, frag_aos, nb, lam, rng))
    return Problem(
        nao=nao, naux=naux, S_abs=np.asfortranarray(S_abs), exch_reachable=exch,
        unique_dense_data=M, L_PQ=L, TAs=TAs, MO_coeff_epsilon=MO_coeff_epsilon,
        AO_coeff_epsilon=AO_coeff_epsilon, name=name,
        meta=dict(n_side=n_side, ao_per_site=ao_per_site, sigma=sigma, lam=lam,
                  cond=cond, seed=seed),
    )


# Named shapes from BASELINE.json / SURVEY.md §8(d) -------------------------

def octane_like(seed: int = SEED, n_frag: int | None = None, matmul=None) -> Problem:
    """nao 202, naux 700, 8 fragments (C8H18 cc-pVDZ / cc-pVDZ-RI, BE3-like)."""
    return make_chain_problem(
        n_units=8, ao_per_unit=24, n_end=5, naux=700, frag_halfwidth=2,
        n_bath=33, seed=seed, n_frag=n_frag, name="octane-like", matmul=matmul,
    )


def c32h66_like(seed: int = SEED, n_frag: int | None = None, matmul=None) -> Problem:
    """nao 778, naux 2716, 32 fragments, nmo <= 220 (C32H66 cc-pVDZ/-RI BE3-like)."""
    return make_chain_problem(
        n_units=32, ao_per_unit=24, n_end=5, naux=2716, frag_halfwidth=2,
        n_bath=100, seed=seed, n_frag=n_frag, name="C32H66-like", matmul=matmul,
    )


# Throughput-sweep points and oversize fragments (BASELINE.json configs[3], configs[4]) ----------------------

def _fit_sigma(r: np.ndarray, naux: int, max_bytes: float, sigma: float, shrink: float = 0.85, floor: float = 0.2) -> float:
    """Largest sigma (from the start value down) whose screened (P|mu nu) fits ``max_bytes`` of host memory: the
    sweep and the oversize cases measure the fragment kernels, not host-side generation."""
    d2 = ((r[:, None, :] - r[None, :, :]) ** 2).sum(-1)
    while True:
        nu = int(np.tril(np.exp(-d2 / (2.0 * sigma * sigma)) >= 1e-10).sum())
        if nu * naux * 8 <= max_bytes or sigma < floor:
            return sigma
        sigma *= shrink


def sweep_point(nmo: int, naux: int, max_bytes: float = 3e9, matmul=None, cond: float = 1e8,
                device_linalg: bool = False) -> Problem:
    """One point of the synthetic sweep (SURVEY 8d "sweep": nmo in 20..400 x naux in 500..8000, nao ~ 4 nmo, one
    fragment): a chain whose interior fragment (``TAs[-1]``) has 3 units (55 % of nmo) + bath tails (45 %)."""
    a = max(2, round(0.55 * nmo / 3))
    n_bath = max(1, nmo - 3 * a)
    n_units = max(5, round(4 * nmo / a))
    sigma = _fit_sigma(chain_centres(n_units, a, 0, np.random.default_rng(1)), naux, max_bytes, 0.94)
    p = make_chain_problem(n_units=n_units, ao_per_unit=a, n_end=0, naux=naux, frag_halfwidth=1, n_bath=n_bath,
                           sigma=sigma, cond=cond, seed=SEED + nmo + naux, n_frag=n_units // 2 + 1, matmul=matmul,
                           name=f"sweep nmo{nmo} naux{naux}", device_linalg=device_linalg)
    p.meta["sigma"] = sigma
    return p


def h2o64_like(n_side: int = 4, naux: int = 9024, n_frag: int = 1, max_bytes: float = 6e9, matmul=None,
               n_sites: int = 4, n_bath: int = 100) -> Problem:
    """(H2O)64-like 3-D lattice (BASELINE.json configs[3]): ``n_side^3`` sites x 58 AOs (cc-pVTZ water), fragments of
    ``n_sites`` sites + ``n_bath`` bath orbitals (nmo 332 by default).  ``n_side=4, naux=9024`` is the full-size case
    (nao 3712); smaller ``n_side`` / ``naux`` keep nmo and shrink the environment so that a CPU oracle finishes."""
    sigma = _fit_sigma(lattice_centres(n_side, 58, np.random.default_rng(7)), naux, max_bytes, 0.6, shrink=0.9, floor=0.1)
    p = make_lattice_problem(n_side=n_side, ao_per_site=58, naux=naux, n_frag=max(n_frag, 1), n_bath=n_bath, sigma=sigma,
                             cond=1e8, seed=SEED + 64, name="(H2O)64-like", matmul=matmul, n_sites=n_sites)
    p.meta["sigma"] = sigma
    return p
