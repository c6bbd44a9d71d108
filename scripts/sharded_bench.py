"""Aux-sharded fragments on N GPUs (torchrun, one rank per GPU): SURVEY 8e "oversize fragments" / acceptance 7.

  torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 scripts/sharded_bench.py --case chain400|h2o64

Cases (synthetic; sigma is reduced until (P|mu nu) fits --max-gb of host memory, like scripts/sweep.py):
  chain400  the nmo = 400, naux = 8000 point of the throughput sweep (BASELINE.json configs[4])
  h2o64     (H2O)64-like 3-D lattice: nao 3712 (64 sites x 58 AOs), naux 9024, corner fragment of 4 sites + bath

Phase A (parity): one fragment through AuxShardedBuilder (slab staging, partial kernels, NCCL reduce-sum onto
rank 0) against the same fragment unsharded on rank 0's GPU, compared on the device; bar 1e-10 max-abs.
Phase B (time): `--frags` fragments sharded, G left on the owners' devices (no D2H), wall clock bracketed by
barrier + device synchronisation on both sides, max over ranks; then the same fragments unsharded on rank 0.
"""
import argparse
import json
import os
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch
import torch.distributed as dist

from quemb_b200 import eri_sparse_DF as q
from quemb_b200 import synthetic as syn
from quemb_b200.sharded import AuxShardedBuilder, FusedAuxShardedBuilder


def gpu_matmul(A, B):
    out = np.empty((A.shape[0], B.shape[1]), order="F")
    Ad = torch.from_numpy(np.ascontiguousarray(A)).cuda()
    for c0 in range(0, B.shape[1], 8192):
        Bd = torch.from_numpy(np.ascontiguousarray(B[:, c0:c0 + 8192])).cuda()
        out[:, c0:c0 + 8192] = (Ad @ Bd).cpu().numpy()
    del Ad
    torch.cuda.empty_cache()
    return out


def fit_sigma(r, naux, max_bytes, sigma):
    while True:
        d2 = ((r[:, None, :] - r[None, :, :]) ** 2).sum(-1)
        nu = int(np.tril(np.exp(-d2 / (2 * sigma * sigma)) >= 1e-10).sum())
        if nu * naux * 8 <= max_bytes or sigma < 0.1:
            return sigma
        sigma *= 0.9


def make_case(case: str, max_bytes: float, nfrag: int):
    rng = np.random.default_rng(7)
    if case == "chain400":
        nmo, naux = 400, 8000
        a = round(0.55 * nmo / 3)
        n_units = round(4 * nmo / a)
        sigma = fit_sigma(syn.chain_centres(n_units, a, 0, rng), naux, max_bytes, 0.94)
        p = syn.make_chain_problem(n_units=n_units, ao_per_unit=a, n_end=0, naux=naux, frag_halfwidth=1, n_bath=nmo - 3 * a,
                                   sigma=sigma, cond=1e8, seed=syn.SEED + 400, matmul=gpu_matmul, name="chain400")
        frags = [n_units // 2 + k for k in range(nfrag)]
    elif case == "h2o64":
        naux = 9024
        sigma = fit_sigma(syn.lattice_centres(4, 58, rng), naux, max_bytes, 0.6)
        p = syn.make_lattice_problem(n_side=4, ao_per_site=58, naux=naux, n_frag=max(nfrag, 1), n_bath=100, sigma=sigma,
                                     cond=1e8, seed=syn.SEED + 64, name="(H2O)64-like", matmul=gpu_matmul, n_sites=4)
        frags = list(range(nfrag))
    elif case == "tiny":
        p = syn.make_chain_problem(n_units=10, ao_per_unit=12, n_end=3, naux=333, frag_halfwidth=1, n_bath=20, sigma=0.7)
        frags = [4 + k for k in range(nfrag)]
        sigma = 0.7
    else:
        raise SystemExit(f"unknown case {case}")
    return p, frags, sigma


class DeviceSinkBuilder(AuxShardedBuilder):
    def _to_host(self, G, npair):  # keep the summed matrix on the device for the on-device comparison
        return G


class PeerMatrixView:
    """Slices of an exported matrix as torch tensors (device copies of the slice only: no second npair^2 buffer)."""

    def __init__(self, ctx, target):
        self.ctx, self.target = ctx, target

    def __getitem__(self, sl):
        from quemb_b200._lib import c_i64, c_p, check, lib

        n = sl.stop - sl.start
        out = torch.empty(n, dtype=torch.float64, device="cuda")
        check(lib().qemb_device_to_host(self.ctx._h, c_p(out.data_ptr()), c_p(self.target.value + 8 * sl.start), c_i64(8 * n)))
        return out


class FusedDeviceSinkBuilder(FusedAuxShardedBuilder):
    def _owner_finish(self, target, npair, to_host):  # mirror on the device; the matrix stays where it is
        from quemb_b200._lib import c_i64, check, lib

        check(lib().qemb_device_mirror_lower(self.ctx._h, target, c_i64(npair), c_i64(npair)))
        self.ctx.synchronize()
        return PeerMatrixView(self.ctx, target) if to_host else None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--case", default="chain400")
    ap.add_argument("--frags", type=int, default=2)
    ap.add_argument("--max-gb", type=float, default=6.0)
    ap.add_argument("--out", default="gpurun_out/sharded.json")
    ap.add_argument("--fused", action="store_true", help="K3 adds its tiles into the owner's matrix over NVLink peer memory")
    args = ap.parse_args()
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()

    def sync():
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()

    t0 = time.time()
    p, frags, sigma = make_case(args.case, args.max_gb * 1e9, args.frags)
    gen_s = time.time() - t0
    shape = (p.naux, p.nao, p.nao)
    TAs = p.TAs
    nmo = TAs[frags[0]].shape[1]
    npair = nmo * (nmo + 1) // 2

    b = (FusedDeviceSinkBuilder if args.fused else DeviceSinkBuilder)(dist, device=local)
    t0 = time.time()
    b.stage(p.unique_dense_data, shape, p.exch_reachable, p.L_PQ, p.S_abs)
    sync()
    stage_s = time.time() - t0

    # ---- phase A: parity of one sharded fragment against the unsharded path on rank 0 --------------------
    kept = {}
    b.build(TAs, p.MO_coeff_epsilon, lambda i, G: kept.__setitem__(i, G), indices=frags[:1])
    sync()
    err = None
    full = metric = screen = None
    if rank == 0:
        ctx = b.ctx
        full = q.SemiSparseSym3DTensor(p.unique_dense_data, shape, p.exch_reachable, ctx=ctx)
        metric = q.GPU_MatrixHandle(p.L_PQ, ctx=ctx)
        full.whiten(metric)
        screen = b.screen
        G1 = torch.empty(npair * npair, dtype=torch.float64, device="cuda")
        q.transform_integral_device(full, TAs[frags[0]], screen, None, p.MO_coeff_epsilon, G1.data_ptr(), npair, ws=b.ws)
        ctx.synchronize()
        Gs = kept[frags[0]]
        err = 0.0
        step = 1 << 28
        for o in range(0, npair * npair, step):  # chunked: no third npair^2 temporary
            e = min(npair * npair, o + step)
            err = max(err, float((Gs[o:e] - G1[o:e]).abs().max()))
        gmax = max(float(G1[o:o + step].abs().max()) for o in range(0, npair * npair, step))
        del G1
    kept.clear()
    sync()

    # ---- phase B: time ---------------------------------------------------------------------------------
    b.build(TAs, p.MO_coeff_epsilon, None, indices=frags)  # warm-up (buffers, plans)
    sync()
    t0 = time.perf_counter()
    b.build(TAs, p.MO_coeff_epsilon, None, indices=frags)
    sync()
    t_sh = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    dist.all_reduce(t_sh, op=dist.ReduceOp.MAX)
    slab_bytes = b.tensor.nonzero_size * 8
    b._G = [None, None]
    kept.clear()
    torch.cuda.empty_cache()
    t_un = None
    if rank == 0:
        G1 = torch.empty(npair * npair, dtype=torch.float64, device="cuda")
        for rep in range(2):
            b.ctx.synchronize()
            t0 = time.perf_counter()
            for i in frags:
                q.transform_integral_device(full, TAs[i], screen, None, p.MO_coeff_epsilon, G1.data_ptr(), npair, ws=b.ws)
            b.ctx.synchronize()
            t_un = time.perf_counter() - t0
        flops = p.naux * float(npair) * (npair + 1) * len(frags)
        line = {"case": args.case, "world": world, "reduce": "fused K3 epilogue -> peer memory (red.add.f64 over NVLink)" if args.fused else "NCCL reduce (async, overlapped with the next fragment)", "nao": p.nao, "naux": p.naux, "nmo": nmo, "npair": npair,
                "n_unique": p.n_unique, "sigma": round(sigma, 3), "fragments": len(frags), "slabs": b.slab,
                "slab_tensor_gb": slab_bytes / 1e9, "full_tensor_gb": p.unique_dense_data.nbytes / 1e9,
                "G_gb": npair * npair * 8 / 1e9, "gen_s": gen_s, "stage_s": stage_s,
                "parity_max_abs_sharded_vs_unsharded": err, "G_max_abs": gmax,
                "sharded_s": float(t_sh.item()), "unsharded_1gpu_s": t_un, "speedup": t_un / float(t_sh.item()),
                "K3_tflops_aggregate_sharded": flops / float(t_sh.item()) / 1e12,
                "reduce_bytes_per_fragment": npair * npair * 8,
                "timing": "wall clock between barrier+cudaDeviceSynchronize pairs, max over ranks; G stays on the device"}
        print(json.dumps(line), flush=True)
        Path(args.out).parent.mkdir(parents=True, exist_ok=True)
        with open(args.out, "a") as f:
            f.write(json.dumps(line) + "\n")
        assert err <= 1e-10, err
    sync()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
