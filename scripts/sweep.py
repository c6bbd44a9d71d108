"""BASELINE.json configs[4]: synthetic fragment-ERI throughput sweep on ONE GPU.

fragment size nmo in {20..400} x naux in {500..8000}, nao = 4 nmo, one fragment per point (SURVEY 8d "sweep").
Per point: device time of K1 / K2a / K3 (CUDA events on the library's compute stream, best of `reps`), the
algorithmic flops of SURVEY 8d, TF/s and the fraction of the FP64 tensor peak (cuBLAS DGEMM 8192^3 measured in
the same process).  (P|mu nu) is whitened once per point (design option O1), so K2b does not appear per fragment;
its one-off time is reported as `whiten_ms`.

    python scripts/sweep.py [--quick] [--out gpurun_out/sweep.json]
"""
import argparse
import json
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch

from quemb_b200 import eri_sparse_DF as q
from quemb_b200 import synthetic as syn


def dgemm_peak(n=8192, iters=6):
    a = torch.randn(n, n, dtype=torch.float64, device="cuda")
    b = torch.randn(n, n, dtype=torch.float64, device="cuda")
    c = a @ b
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        c = a @ b
    e1.record()
    torch.cuda.synchronize()
    return iters * 2 * n ** 3 / (e0.elapsed_time(e1) * 1e-3) / 1e12


def gpu_matmul(A, B):
    out = np.empty((A.shape[0], B.shape[1]), order="F")
    Ad = torch.from_numpy(np.ascontiguousarray(A)).cuda()
    for c0 in range(0, B.shape[1], 16384):
        Bd = torch.from_numpy(np.ascontiguousarray(B[:, c0:c0 + 16384])).cuda()
        out[:, c0:c0 + 16384] = (Ad @ Bd).cpu().numpy()
    return out


def make_point(nmo: int, naux: int, max_bytes: float):
    """The shared generator (quemb_b200.synthetic.sweep_point) with the GPU-accelerated L @ Y."""
    p = syn.sweep_point(nmo, naux, max_bytes, matmul=gpu_matmul)
    return p, p.meta["sigma"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--batch", type=int, default=0,
                    help="additionally time a batch of that many fragments of the point's size through 1/2/4/8 concurrent "
                         "lanes (FragmentERIBuilder.build_resident): aggregate TF/s of small fragments")
    ap.add_argument("--low", action="store_true", help="only the low end: nmo {20, 50, 100} x naux {500, 2000}")
    ap.add_argument("--out", default="gpurun_out/sweep.json")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--max-gb", type=float, default=3.0)
    args = ap.parse_args()
    nmos = [20, 100, 400] if args.quick else [20, 50, 100, 150, 200, 300, 400]
    nauxs = [500, 8000] if args.quick else [500, 1000, 2000, 4000, 8000]
    if args.low:
        nmos, nauxs = [20, 50, 100], [500, 2000]
    peak = dgemm_peak()
    ctx = q.default_context(0)
    ctx.set_timing(True)
    ws = q.default_workspace(ctx)
    rows = []
    print(f"cuBLAS DGEMM 8192^3: {peak:.2f} TF/s", flush=True)
    for naux in nauxs:
        for nmo in nmos:
            t0 = time.time()
            p, sigma = make_point(nmo, naux, args.max_gb * 1e9)
            TA = p.TAs[-1]  # the interior fragment (full 3 units + bath)
            n = TA.shape[1]
            npair = n * (n + 1) // 2
            T = q.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable, ctx=ctx)
            L = q.GPU_MatrixHandle(p.L_PQ, ctx=ctx)
            tw = time.time()
            T.whiten(L)
            ctx.synchronize()
            whiten_ms = (time.time() - tw) * 1e3
            S = q._ScreenHandle(p.S_abs, ctx)
            G = torch.empty(npair * npair, dtype=torch.float64, device="cuda")
            best = None
            for _ in range(args.reps + 1):
                q.transform_integral_device(T, TA, S, None, p.MO_coeff_epsilon, G.data_ptr(), npair, ws=ws)
                ctx.synchronize()
                ms = ws.times_ms()
                best = ms if best is None else {k: min(best[k], ms[k]) for k in ms}
            wk = ws.work()
            row = {"nmo": n, "naux": naux, "nao": p.nao, "npair": npair, "n_unique": p.n_unique, "sigma": round(sigma, 3),
                   "whiten_ms": whiten_ms, "gen_s": round(time.time() - t0, 1)}
            tot_f, tot_ms = 0.0, 0.0
            for k, fk, bk in (("K1", "F_K1", "B_K1"), ("K2a", "F_K2a", "B_K2a"), ("K3", "F_K3", "B_K3")):
                t = best[k] * 1e-3
                row[k] = {"ms": best[k], "tflops": wk[fk] / t / 1e12 if t > 0 else None,
                          "frac_fp64_peak": wk[fk] / t / 1e12 / peak if t > 0 else None,
                          "gbs": wk[bk] / t / 1e9 if t > 0 else None}
                tot_f += wk[fk]
                tot_ms += best[k]
            # J/K consumer on the resident matrix (HBM bound: 8 npair^2 bytes read once)
            rng = np.random.default_rng(0)
            D = rng.standard_normal((n, n))
            D = D + D.T
            jk_ms = None
            for _ in range(2):
                ws.jk(D)
                ms_jk, nbytes = ws.jk_time()
                jk_ms = ms_jk if jk_ms is None else min(jk_ms, ms_jk)
            row["JK"] = {"ms": jk_ms, "gbs": nbytes / (jk_ms * 1e-3) / 1e9 if jk_ms else None}
            row["total_ms"] = tot_ms + best["screen"]
            row["total_tflops"] = tot_f / (row["total_ms"] * 1e-3) / 1e12
            if args.batch:
                from quemb_b200.driver import FragmentERIBuilder

                bb = FragmentERIBuilder(0)
                bb.stage(None, None, None, L, S, whiten=False, tensor=T)  # T is whitened already
                same = [i for i, t in enumerate(p.TAs) if t.shape[1] == n] or [len(p.TAs) - 1]
                idx = [same[k % len(same)] for k in range(args.batch)]
                row["batch"] = {"fragments": args.batch, "flops_per_fragment": tot_f}
                for nt in (1, 2, 4, 8):
                    bb.build_resident(p.TAs, p.MO_coeff_epsilon, indices=idx, n_threads=nt)  # warm-up (lanes, plans)
                    ctx.synchronize()
                    bestb = None
                    for _ in range(3):
                        ctx.mark(8)
                        bb.build_resident(p.TAs, p.MO_coeff_epsilon, indices=idx, n_threads=nt)
                        ctx.mark(9)
                        ctx.synchronize()
                        msb = ctx.mark_elapsed_ms(8, 9)
                        bestb = msb if bestb is None else min(bestb, msb)
                    tf = tot_f * args.batch / (bestb * 1e-3) / 1e12
                    row["batch"][f"lanes_{nt}"] = {"ms": bestb, "tflops": tf, "frac_fp64_peak": tf / peak}
                del bb
            rows.append(row)
            print(json.dumps(row), flush=True)
            del T, L, S, G, p
            torch.cuda.empty_cache()
    Path(args.out).parent.mkdir(parents=True, exist_ok=True)
    Path(args.out).write_text(json.dumps({"fp64_peak_tflops_cublas": peak, "points": rows}, indent=1))


if __name__ == "__main__":
    main()
