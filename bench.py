#!/usr/bin/env python
"""bench.py -- all-fragment ERI build time of the sparse-DF path (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W          B200 arm (this repo's CUDA path)
  python bench.py --impl reference --gpus N ...           the reference's own CPU path on the host cores

A "step" is ONE all-fragment build of the workload molecule: (re)apply the Coulomb metric to the
resident (P|mu nu), then screen -> K1 -> K2a -> K3 for each of its fragments.  Workload at every N:
the C32H66-like synthetic alkane (BASELINE.json configs[2], the configuration the metric is quoted on:
nao 778, naux 2716, 32 BE3-like fragments); total work is fixed, fragments are partitioned over the N
ranks (strong scaling, no data-path collective).

  value  seconds per build with all inputs resident in HBM and G left on the device (kernel path)
  e2e    seconds per build through the public host API: (P|mu nu), L, S_abs and every TA start in host
         memory, every fragment's G ends in (pinned) host memory; host<->device copies inside the timing
Both are device-timed with CUDA events (start mark on the idle compute stream, end mark after the last
kernel / copy), max over ranks.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "all_fragment_eri_build_time"
UNIT = "s"


# ---------------------------------------------------------------------------------------------------
# workload
# ---------------------------------------------------------------------------------------------------
def make_workload(name: str, use_gpu_matmul: bool):
    from quemb_b200 import synthetic as syn

    mm = None
    if use_gpu_matmul:
        import torch

        def mm(A, B):  # synthetic-data generation only (L @ Y), not part of any timed region
            out = np.empty((A.shape[0], B.shape[1]), order="F")
            Ad = torch.from_numpy(np.ascontiguousarray(A)).cuda()
            step = 16384
            for c0 in range(0, B.shape[1], step):
                Bd = torch.from_numpy(np.ascontiguousarray(B[:, c0:c0 + step])).cuda()
                out[:, c0:c0 + step] = (Ad @ Bd).cpu().numpy()
            return out

    if name == "c32h66":
        p = syn.c32h66_like(matmul=mm)
    elif name == "octane":
        p = syn.octane_like(matmul=mm)
    elif name == "tiny":  # CI / CPU smoke of the harness itself
        p = syn.make_chain_problem(n_units=8, ao_per_unit=8, n_end=2, naux=160, frag_halfwidth=1, n_bath=8,
                                   sigma=0.6, name="tiny", matmul=mm)
    else:
        raise SystemExit(f"unknown workload {name}")
    return p


def workload_desc(p) -> str:
    nmos = [t.shape[1] for t in p.TAs]
    return (f"{p.name} synthetic chain: nao {p.nao}, naux {p.naux}, n_unique {p.n_unique} screened AO pairs "
            f"(AO eps {p.AO_coeff_epsilon:g}), {len(p.TAs)} fragments, nmo {min(nmos)}-{max(nmos)}, "
            f"MO eps {p.MO_coeff_epsilon:g}, metric cond {p.meta.get('cond', 0):g}, seed {p.meta.get('seed')}")


def reference_executed_flops(p, TA, A) -> float:
    """What the reference's CPU path executes for one fragment: K1 and K2a
    axpys, trsm, and a FULL npair x npair GEMM (cpp:620 is not a syrk)."""
    nmo = TA.shape[1]
    npair = nmo * (nmo + 1) // 2
    rlen = np.array([len(x) for x in p.exch_reachable], dtype=np.float64)
    f1 = 2.0 * p.naux * sum(float(rlen[np.asarray(a, dtype=np.int64)].sum()) for a in A if len(a))
    f2 = 2.0 * p.naux * sum(len(A[a]) * (nmo - a) for a in range(nmo))
    return f1 + f2 + float(p.naux) ** 2 * npair + 2.0 * p.naux * float(npair) ** 2


# ---------------------------------------------------------------------------------------------------
# clocks sampler
# ---------------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.path = tempfile.NamedTemporaryFile(prefix="clocks_", suffix=".csv", delete=False).name
        self.proc = None
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200", "-i",
                 str(gpu_index)], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except OSError:
            self.proc = None

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in open(self.path):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1])); power.append(float(f[2]))
            except ValueError:
                continue
            for n, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        os.unlink(self.path)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        busy = [s for s, w in zip(sm, power) if w > 0.5 * max(power)] or sm
        return {"sm_mhz": float(np.median(busy)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons),
                "power_w_max": float(max(power)), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------
# reference arm: the reference's own CPU implementation on the host cores
# ---------------------------------------------------------------------------------------------------
def reference_sample_step(ref, kind, p, T_ref, frag: int):
    """One bounded sample: the reference's transform_integral on ONE fragment."""
    from oracle import oracle

    TA = p.TAs[frag]
    t = time.perf_counter()
    if ref is not None:
        G = ref.transform_integral(T_ref, TA, p.S_abs, p.L_PQ, p.MO_coeff_epsilon)
    else:
        G = oracle.transform_integral(p.unique_dense_data, p.exch_reachable, TA, p.S_abs, p.L_PQ, p.MO_coeff_epsilon)
    dt = time.perf_counter() - t
    return dt, G


def reference_setup(p):
    from oracle import oracle

    ref = oracle.reference_module()
    kind = "reference" if ref is not None else "port"
    T_ref = None
    if ref is not None:
        T_ref = ref.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao),
                                          [x.tolist() for x in p.exch_reachable])
    works = []
    for TA in p.TAs:
        A = oracle.get_AO_per_MO(TA, p.S_abs, p.MO_coeff_epsilon)
        works.append(reference_executed_flops(p, TA, A))
    return ref, kind, T_ref, works


def pick_sample_fragment(p, works, budget_flops: float) -> int:
    """Largest fragment whose executed work stays under the budget (else the smallest one)."""
    ok = [i for i, w in enumerate(works) if w <= budget_flops]
    return max(ok, key=lambda i: works[i]) if ok else int(np.argmin(works))


def run_reference(args, rank: int, world: int) -> None:
    if rank != 0:
        return
    p = make_workload(args.workload, use_gpu_matmul=False)
    ref, kind, T_ref, works = reference_setup(p)
    cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    # ~8 s per step at ~150 GF/s sustained on a 16-core host
    frag = pick_sample_fragment(p, works, 1.2e12)
    scale = sum(works) / works[frag]
    for _ in range(args.warmup):
        reference_sample_step(ref, kind, p, T_ref, frag)
    ts = [reference_sample_step(ref, kind, p, T_ref, frag)[0] for _ in range(args.steps)]
    t_sample = float(np.mean(ts))
    value = t_sample * scale
    sample = (f"fragment {frag} (nmo {p.TAs[frag].shape[1]}) of {len(p.TAs)} per step, {t_sample:.2f} s, extrapolated to "
              f"all fragments by the reference's executed flops (x{scale:.1f}); "
              f"{'oracle/_ref = unmodified reference extension' if kind == 'reference' else 'oracle C port'}, OpenMP over {cores} threads")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": value * 1e3, "higher_is_better": False, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_desc(p)},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------------
# B200 arm
# ---------------------------------------------------------------------------------------------------
def fp64_peak_tflops(torch, n=8192, secs=1.5):
    """cuBLAS DGEMM n^3: burst (best single) and sustained (back to back for `secs`)."""
    a = torch.randn(n, n, dtype=torch.float64, device="cuda")
    b = torch.randn(n, n, dtype=torch.float64, device="cuda")
    c = a @ b
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 0.0
    for _ in range(5):
        e0.record(); c = a @ b; e1.record(); torch.cuda.synchronize()
        best = max(best, 2 * n ** 3 / (e0.elapsed_time(e1) * 1e-3) / 1e12)
    t0 = time.perf_counter(); reps = 0
    e0.record()
    while time.perf_counter() - t0 < secs:
        for _ in range(4):
            c = a @ b
        reps += 4
        torch.cuda.synchronize()
    e1.record(); torch.cuda.synchronize()
    sustained = reps * 2 * n ** 3 / (e0.elapsed_time(e1) * 1e-3) / 1e12
    del a, b, c
    torch.cuda.empty_cache()
    return best, sustained


def run_b200(args, rank: int, world: int, local_rank: int) -> None:
    import torch

    from quemb_b200 import _lib
    from quemb_b200 import eri_sparse_DF as q
    from quemb_b200.driver import FragmentERIBuilder, fragment_cost, partition_lpt

    if not torch.cuda.is_available() or _lib.device_count() == 0:
        raise SystemExit("bench.py: no CUDA device -- the B200 arm has no CPU fallback (use --impl reference)")
    dist = None
    if world > 1:
        import torch.distributed as dist

        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    torch.cuda.set_device(local_rank)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x: float) -> float:
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    peak_burst, peak_sust = fp64_peak_tflops(torch) if rank == 0 else (0.0, 0.0)

    p = make_workload(args.workload, use_gpu_matmul=True)
    nfrag = len(p.TAs)
    costs = [fragment_cost(p.naux, t.shape[1]) for t in p.TAs]
    mine = partition_lpt(costs, world)[rank]
    max_nmo = max(t.shape[1] for t in p.TAs)
    max_npair = max_nmo * (max_nmo + 1) // 2

    b = FragmentERIBuilder(local_rank)
    ctx = b.ctx
    ctx.set_k3_int8_slices(args.k3_int8)
    # raw (un-whitened) device copy kept aside so that every step redoes the metric application
    t0 = time.perf_counter()
    raw = q.SemiSparseSym3DTensor(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable, ctx=ctx)
    work_t = q.SemiSparseSym3DTensor((p.naux, p.nao, p.nao), p.exch_reachable, ctx=ctx)
    b.stage(None, None, None, p.L_PQ, p.S_abs, whiten=False, tensor=work_t)
    staging_s = time.perf_counter() - t0
    Gdev = torch.empty(max_npair * max_npair, dtype=torch.float64, device="cuda")
    b.reserve_host(max_nmo)
    ctx.set_timing(True)

    ktimes = {k: 0.0 for k in ("screen", "K1", "K2a", "K2b", "K3")}
    kwork = {}
    klaunch = {"K3": 0}

    def device_step(collect: bool):
        work_t.copy_data_from(raw)
        if dist is None:
            work_t.whiten(b.metric)
        else:  # every rank needs the whole whitened tensor: 1/N of the solve each + NCCL broadcasts over NVLink
            work_t.whiten_distributed(b.metric, dist)
        for i in mine:
            b.build_device(p.TAs, p.MO_coeff_epsilon, Gdev.data_ptr(), indices=[i])
            if collect:  # per-kernel CUDA-event times of this fragment (waits for it: not used in timed steps)
                ms = b.ws.times_ms(); wk = b.ws.work()
                for k in ktimes:
                    ktimes[k] += ms[k]
                for k, v in wk.items():
                    kwork[k] = kwork.get(k, 0.0) + v
                klaunch["K3"] += 1

    # ---- value: device-resident build --------------------------------------------------------------
    for _ in range(args.warmup):
        device_step(False)
    ctx.synchronize(); barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    l0 = ctx.launch_count
    ctx.mark(0, "compute")
    for _ in range(args.steps):
        device_step(False)
    ctx.mark(1, "compute")
    ctx.synchronize(); barrier()
    launches = ctx.launch_count - l0
    t_dev = max_over_ranks(ctx.mark_elapsed_ms(0, 1) * 1e-3) / args.steps
    # one extra untimed pass that collects per-kernel event times (each fragment waited for)
    device_step(True)
    ctx.synchronize()
    # the consumer right behind the path, informational: J/K of one density from the (ij|kl) the
    # last fragment left on the device (HBM bound: 8 npair^2 bytes read once)
    consumer = None
    if rank == 0 and mine:
        n_last = p.TAs[mine[-1]].shape[1]
        Dm = np.random.default_rng(1).standard_normal((n_last, n_last))
        Dm = Dm + Dm.T
        best, nb = None, 0.0
        for _ in range(4):
            b.ws.jk(Dm)
            ms_jk, nb = b.ws.jk_time()
            best = ms_jk if best is None else min(best, ms_jk)
        consumer = {"kernel": "jk_stream_kernel + jk_reduce_kernel (J and K of one density from the resident (ij|kl))",
                    "fragment_nmo": int(n_last), "ms": best, "gbs": nb / best / 1e6}

    # ---- e2e: host buffers in, pinned host buffers out ------------------------------------------------
    d2h = sum(8 * (t.shape[1] * (t.shape[1] + 1) // 2) ** 2 for t in (p.TAs[i] for i in mine))
    h2d = (p.unique_dense_data.nbytes + p.L_PQ.nbytes + p.S_abs.nbytes + sum(p.TAs[i].nbytes for i in mine))
    checks = []

    def sink(i, G):
        checks.append(float(G[0, 0]))  # the step's result is read on the host

    def e2e_step():
        # everything a caller holding HOST arrays pays: upload (P|mu nu), L, S_abs; whiten; all fragments out
        work_t.upload(p.unique_dense_data)
        b.stage(None, None, None, p.L_PQ, p.S_abs, whiten=True, tensor=work_t)
        b.build(p.TAs, p.MO_coeff_epsilon, sink, indices=mine)

    e2e_steps = max(1, min(args.steps, 3))
    e2e_step()
    ctx.synchronize(); barrier()
    xfer0 = ctx.transfer_bytes
    ctx.mark(2, "compute")
    tw = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    ctx.mark(3, "copy")
    ctx.synchronize(); barrier()
    wall_e2e = (time.perf_counter() - tw) / e2e_steps
    xfer1 = ctx.transfer_bytes
    # bytes the copy engines actually moved per step (G travels as its lower triangle; counted by the library)
    h2d, d2h_moved = (xfer1[0] - xfer0[0]) / e2e_steps, (xfer1[1] - xfer0[1]) / e2e_steps
    d2h_result = d2h
    t_e2e = max_over_ranks(ctx.mark_elapsed_ms(2, 3) * 1e-3) / e2e_steps
    clocks = sampler.stop() if sampler else None
    h2d_all, d2h_all = sum_over_ranks(float(h2d)), sum_over_ranks(float(d2h_moved))
    d2h_result_all = sum_over_ranks(float(d2h_result))
    launches_all = int(sum_over_ranks(float(launches)))

    # ---- cpu baseline (rank 0, N = 1 only): bounded sample of the reference's CPU path ------------------
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        ref, kind, T_ref, works = reference_setup(p)
        cores = len(os.sched_getaffinity(0))
        frag = pick_sample_fragment(p, works, 1.2e12)
        dt, _ = reference_sample_step(ref, kind, p, T_ref, frag)
        dt2, G_ref = reference_sample_step(ref, kind, p, T_ref, frag)
        dt = min(dt, dt2)
        scale = sum(works) / works[frag]
        # the same fragment through the B200 path (whitened tensor as timed above): full-size parity, untimed
        got = {}
        b.build(p.TAs, p.MO_coeff_epsilon, lambda i, G: got.__setitem__(i, float(np.abs(G - G_ref).max())), indices=[frag])
        cpu_baseline = {
            "value": dt * scale, "unit": UNIT, "cores": cores, "kind": kind,
            "sample": (f"fragment {frag} (nmo {p.TAs[frag].shape[1]}): {dt:.2f} s (best of 2), extrapolated to all "
                       f"{nfrag} fragments by the reference's executed flops (x{scale:.1f}); OpenMP over {cores} threads"),
            "parity_max_abs": got.get(frag),
            "parity_note": f"max |(ij|kl)_B200 - (ij|kl)_reference| over the full {G_ref.shape[0]}^2 matrix of fragment {frag}; bar 1e-10",
        }

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (K3, FP64 tensor pipe) ------------------------------------------
    n_k3 = max(klaunch["K3"], 1)
    k3_ms = ktimes["K3"] / n_k3
    k3_flops = kwork.get("F_K3", 0.0) / n_k3
    achieved = k3_flops / (k3_ms * 1e-3) / 1e12 if k3_ms > 0 else 0.0
    traffic = None
    tj = ROOT / "profiles" / "k3_traffic.json"
    if tj.exists():
        try:
            traffic = json.loads(tj.read_text()).get("dram_bytes_per_launch")
        except Exception:  # noqa: BLE001
            traffic = None
    per_kernel = {}
    for k, fk, bk in (("K1", "F_K1", "B_K1"), ("K2a", "F_K2a", "B_K2a"), ("K3", "F_K3", "B_K3")):
        if ktimes[k] > 0:
            per_kernel[k] = {"ms_total": ktimes[k], "tflops": kwork.get(fk, 0) / (ktimes[k] * 1e-3) / 1e12,
                             "gbs": kwork.get(bk, 0) / (ktimes[k] * 1e-3) / 1e9}
    peaks = {}
    mp = ROOT / "MEASURED_PEAKS.json"
    if mp.exists():
        peaks = json.loads(mp.read_text())
    if consumer and peaks.get("hbm_gbs"):
        consumer["hbm_frac"] = consumer["gbs"] / peaks["hbm_gbs"]
    line = {
        "metric": METRIC, "value": t_dev, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": t_dev * 1e3, "higher_is_better": False, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload_desc(p),
                   "l2": "inputs larger than L2: whitened (P|mu nu) 2.4 GB, X 0.4-0.55 GB and G 1.1-5.2 GB per fragment",
                   "partition": (f"fragments LPT over {world} rank(s), no collective on the fragment path; the one-off metric "
                                 "application is split over the ranks by columns and exchanged with NCCL broadcasts"
                                 if world > 1 else "1 rank"),
                   "fragments_rank0": len(mine),
                   "k3": (f"opt-in int8-slice K3, {args.k3_int8} slices (roofline.achieved counts the FP64-equivalent flops)"
                          if args.k3_int8 else "FP64 DMMA (default)")},
        "fragments_per_s": nfrag / t_dev,
        "roofline": {"bound": "tensor", "kernel": "gram_kernel (K3: G = X^T X, FP64 DMMA.8x8x4)",
                     "achieved": achieved, "peak": peak_sust, "unit": "TFLOP/s",
                     "frac": achieved / peak_sust if peak_sust else None, "traffic": traffic,
                     "peak_source": "cuBLAS FP64 DGEMM 8192^3 measured in this run, sustained "
                                    f"(burst {peak_burst:.1f}); MEASURED_PEAKS.json has no FP64 entry",
                     "algorithmic_flops_per_launch": k3_flops, "avg_launch_ms": k3_ms, "launches_measured": n_k3},
        "per_kernel": per_kernel,
        "hbm_peak_gbs_measured": peaks.get("hbm_gbs"),
        "staging_s": staging_s,
        "consumer": consumer,
        "cpu_baseline": cpu_baseline,
        "e2e": {"value": t_e2e, "unit": UNIT, "h2d_bytes_per_step": int(h2d_all), "d2h_bytes_per_step": int(d2h_all),
                "result_bytes_per_step": int(d2h_result_all), "wall_s": wall_e2e, "steps": e2e_steps,
                "note": "includes upload of (P|mu nu)+L+S_abs from pageable host memory, metric application, all "
                        "fragments, every full npair x npair G delivered in pinned host memory (lower triangle over "
                        "PCIe, upper triangle rebuilt by the library's host threads inside the timed region)"},
        "gpu_launches": launches_all,
        "clocks": clocks,
    }
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c32h66", choices=["c32h66", "octane", "tiny"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--k3-int8", type=int, default=0, choices=[0, 5, 6, 7, 8],
                    help="opt-in K3 on the int8 tensor cores with that many slices (default 0: FP64 tensor path)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_b200(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
