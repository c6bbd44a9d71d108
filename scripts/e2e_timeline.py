"""Where the wall time of one end-to-end build goes, per rank (host view), at N = 1 / 2 / 4 / 8.

  python scripts/e2e_timeline.py [--modes triangle:replicated,triangle:sharded,full:sharded] [--reps 3]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 scripts/e2e_timeline.py ...

Per rank and repetition: upload (host -> device of (P|mu nu)), stage (L, S_abs, metric application incl. the NCCL
exchange), build (all of the rank's fragments, every full (ij|kl) in pinned host memory), split into the time the
host spent enqueueing transforms, waiting for outputs and inside the sink; bytes over PCIe both ways; the host's
DRAM bandwidth with the rank's own pool threads.  Rank 0 gathers everything and appends one JSON line to --out.
"""
import argparse
import json
import os
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np

import bench
from quemb_b200 import _lib
from quemb_b200 import eri_sparse_DF as q
from quemb_b200.driver import DistributedQueue, FragmentERIBuilder, fragment_cost, partition_lpt


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--modes", default="triangle:sharded,full:sharded",
                    help="comma list of <output>:<upload>[:dynamic]; output = triangle | full | auto (per-rank balance of "
                         "PCIe against the host mirror, qemb_ctx_autotune_output), upload = replicated | sharded, "
                         "dynamic = fragments handed out across ranks from a shared counter instead of a static partition")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--buffers", type=int, default=3)
    ap.add_argument("--out", default="gpurun_out/e2e_timeline.jsonl")
    args = ap.parse_args()
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    import torch

    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    p = bench.make_workload("c32h66", use_gpu_matmul=True)
    costs = [fragment_cost(p.naux, t.shape[1]) for t in p.TAs]
    mine = partition_lpt(costs, world)[rank]
    b = FragmentERIBuilder(local, n_host_buffers=args.buffers)
    ctx = b.ctx
    T = q.SemiSparseSym3DTensor((p.naux, p.nao, p.nao), p.exch_reachable, ctx=ctx)
    b.reserve_host(max(t.shape[1] for t in p.TAs))
    hw, hc = q.host_bandwidth(1 << 30)

    log = {}
    orig_t, orig_w = q.transform_integral, b.ws.wait_output

    def timed(name, fn):
        def f(*a, **k):
            t = time.perf_counter()
            r = fn(*a, **k)
            log[name] = log.get(name, 0.0) + time.perf_counter() - t
            return r
        return f

    q.transform_integral = timed("enqueue_s", orig_t)
    b.ws.wait_output = timed("wait_s", orig_w)
    chk = []

    def sink(i, G):
        t = time.perf_counter()
        chk.append(float(G[0, 0]) + float(G[-1, 0]) + float(G[0, -1]))
        log["sink_s"] = log.get("sink_s", 0.0) + time.perf_counter() - t

    for mode in args.modes.split(","):
        parts = mode.split(":")
        output, upload, dynamic = parts[0], parts[1], len(parts) > 2 and parts[2] == "dynamic" and dist is not None
        barrier()
        tuned = None
        if output == "auto":
            tuned = b.tune_output()
        else:  # "triangle", "full", or "triangle@<fraction of the columns that cross PCIe whole>"
            fields = output.split("@")  # kind[@fraction[@group MB]]
            kind, frac = fields[0], (fields[1] if len(fields) > 1 else "")
            ctx.set_output_mode(kind == "triangle")
            ctx.set_full_column_fraction(float(frac) if frac else 0.0)
            ctx.set_output_group_bytes(int(float(fields[2]) * (1 << 20)) if len(fields) > 2 else 256 << 20)
        order_all = sorted(range(len(p.TAs)), key=lambda i: (-costs[i], i))
        rows = []
        for rep in range(args.reps + 1):
            log.clear()
            barrier()
            x0 = ctx.transfer_bytes
            t0 = time.perf_counter()
            if world == 1 or upload == "replicated":
                T.upload(p.unique_dense_data)
                ctx.synchronize()
                t1 = time.perf_counter()
                b.stage(None, None, None, p.L_PQ, p.S_abs, whiten=False, tensor=T)
                if dist is None:
                    T.whiten(b.metric)
                else:
                    T.whiten_distributed(b.metric, dist)
            else:
                t1 = t0
                b.stage_distributed(p.unique_dense_data, p.L_PQ, p.S_abs, dist, tensor=T)
            ctx.synchronize()
            t2 = time.perf_counter()
            done = mine
            if dynamic:
                done = b.build_dynamic(p.TAs, p.MO_coeff_epsilon, sink, DistributedQueue(dist, order_all, f"{mode}/{rep}"))
            else:
                b.build(p.TAs, p.MO_coeff_epsilon, sink, indices=mine)
            ctx.synchronize()
            t3 = time.perf_counter()
            barrier()
            t4 = time.perf_counter()
            x1 = ctx.transfer_bytes
            if rep == 0:
                continue  # warm-up
            res = sum(8 * (p.TAs[i].shape[1] * (p.TAs[i].shape[1] + 1) // 2) ** 2 for i in done)
            rows.append({"rep": rep, "fragments_done": len(done), "upload_s": t1 - t0, "stage_s": t2 - t1, "build_s": t3 - t2, "total_s": t3 - t0,
                         "all_ranks_done_s": t4 - t0, **{k: v for k, v in log.items()},
                         "h2d_gb": (x1[0] - x0[0]) / 1e9, "d2h_gb": (x1[1] - x0[1]) / 1e9, "result_gb": res / 1e9,
                         "delivery_gbs": res / (t3 - t2) / 1e9})
        mean = {k: float(np.mean([r[k] for r in rows])) for k in rows[0] if k != "rep"}
        mine_rec = {"rank": rank, "fragments": len(mine), "autotune": tuned, "host_threads": bench.q_host_threads(_lib),
                    "host_write_gbs_rank_threads": hw, "host_copy_gbs_rank_threads": hc, "mean": mean, "reps": rows}
        allrec = [mine_rec]
        if dist is not None:
            allrec = [None] * world
            dist.all_gather_object(allrec, mine_rec)
        if rank == 0:
            line = {"n_gpus": world, "output": output, "upload": upload, "dynamic_queue": bool(dynamic), "buffers": args.buffers,
                    "full_column_fraction_env": os.environ.get("QEMB_FULL_COLUMN_FRACTION"),
                    "cores": len(os.sched_getaffinity(0)),
                    "e2e_s_max_over_ranks": max(r["mean"]["total_s"] for r in allrec),
                    "all_ranks_done_s": max(r["mean"]["all_ranks_done_s"] for r in allrec),
                    "result_gb_total": sum(r["mean"]["result_gb"] for r in allrec),
                    "ranks": allrec}
            line["aggregate_delivery_gbs"] = line["result_gb_total"] / line["e2e_s_max_over_ranks"]
            Path(args.out).parent.mkdir(parents=True, exist_ok=True)
            with open(args.out, "a") as f:
                f.write(json.dumps(line) + "\n")
            brief = {k: v for k, v in line.items() if k != "ranks"}
            brief["rank0_mean"] = allrec[0]["mean"]
            print(json.dumps(brief), flush=True)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
