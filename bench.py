#!/usr/bin/env python
"""bench.py -- all-fragment ERI build time of the sparse-DF path (BASELINE.json metric).

  python bench.py --gpus N --steps K --warmup W          B200 arm (this repo's CUDA path)
  python bench.py --impl reference --gpus N ...           the reference's own CPU path on the host cores

A "step" is ONE all-fragment build of the workload molecule: (re)apply the Coulomb metric to the
resident (P|mu nu), then screen -> K1 -> K2a -> K3 for each of its fragments.  Workload at every N:
the C32H66-like synthetic alkane (BASELINE.json configs[2], the configuration the metric is quoted on:
nao 778, naux 2716, 32 BE3-like fragments); total work is fixed, fragments are partitioned over the N
ranks (strong scaling, no data-path collective).

  value         seconds per build with all inputs resident in HBM and G left on the device (kernel path)
  e2e           seconds per build through the public host API: (P|mu nu), L, S_abs and every TA start in host
                memory, every fragment's FULL npair x npair G ends in (pinned) host memory; copies inside the timing
  e2e_consumer  the same build with the fragment-initialisation consumers (get_veff + update_ebe_hf two-electron
                sums, mbe.py:1146-1170) evaluated where (ij|kl) lives: only nmo x nmo matrices leave the device
  e2e_file      a bounded sample of the delivery-to-file path (background writer fed from the pinned ring)
  sharded       (N > 1) one oversize fragment with the auxiliary index split over the ranks + NCCL reduce-sum
All device-timed with CUDA events (start mark on the idle compute stream, end mark after the last kernel / copy),
max over ranks.
"""
from __future__ import annotations

import os
import sys

if "--impl" in sys.argv and "reference" in sys.argv:
    # torch.distributed.run exports OMP_NUM_THREADS=1 for nproc-per-node > 1; the reference arm is ONE process that
    # owns all host cores.  Must happen before numpy / OpenBLAS / libgomp read the variable.
    _cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
    os.environ["OMP_NUM_THREADS"] = str(_cores)
    os.environ.pop("OMP_THREAD_LIMIT", None)

import argparse  # noqa: E402
import ctypes  # noqa: E402
import json  # noqa: E402
import shutil  # noqa: E402
import subprocess  # noqa: E402
import tempfile  # noqa: E402
import time  # noqa: E402
from pathlib import Path  # noqa: E402

import numpy as np  # noqa: E402

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

METRIC = "all_fragment_eri_build_time"
UNIT = "s"


# ---------------------------------------------------------------------------------------------------
# workload
# ---------------------------------------------------------------------------------------------------
def gpu_matmul(A, B):
    """Synthetic-data generation only (L @ Y), never inside a timed region."""
    import torch

    out = np.empty((A.shape[0], B.shape[1]), order="F")
    Ad = torch.from_numpy(np.ascontiguousarray(A)).cuda()
    step = 16384
    for c0 in range(0, B.shape[1], step):
        Bd = torch.from_numpy(np.ascontiguousarray(B[:, c0:c0 + step])).cuda()
        out[:, c0:c0 + step] = (Ad @ Bd).cpu().numpy()
    del Ad
    torch.cuda.empty_cache()
    return out


def make_workload(name: str, use_gpu_matmul: bool):
    from quemb_b200 import synthetic as syn

    mm = gpu_matmul if use_gpu_matmul else None
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


def config_dict(p, n_gpus: int) -> dict:
    """The SAME dict in both arms (the driver compares them): what the workload is, why L2 cannot help between
    timed iterations, and how the fragments are spread over the GPUs."""
    nmos = [t.shape[1] for t in p.TAs]
    gmin, gmax = (8 * (n * (n + 1) // 2) ** 2 / 1e9 for n in (min(nmos), max(nmos)))
    return {
        "workload": workload_desc(p),
        "l2": (f"inputs larger than L2 (126 MB): (P|mu nu) {p.unique_dense_data.nbytes / 1e9:.1f} GB and "
               f"{gmin:.1f}-{gmax:.1f} GB of (ij|kl) per fragment stream through HBM every step"),
        "partition": (f"fragments LPT over {n_gpus} GPU(s), strong scaling, no collective on the fragment path"),
    }


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
def host_cores() -> int:
    return len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)


class OpenMP:
    """The libgomp instance the reference extension is linked against: thread count per calling thread."""

    def __init__(self):
        try:
            self.lib = ctypes.CDLL("libgomp.so.1")
            self.lib.omp_get_max_threads.restype = ctypes.c_int
        except OSError:
            self.lib = None

    def set(self, n: int) -> None:
        if self.lib is not None:
            self.lib.omp_set_num_threads(ctypes.c_int(int(n)))

    def get(self) -> int | None:
        return int(self.lib.omp_get_max_threads()) if self.lib is not None else None


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


def extrapolate_all_fragments(works, samples: dict[int, float]) -> float:
    """Seconds for ALL fragments from per-fragment samples {fragment: seconds}: every fragment is charged its own
    executed flops at a rate interpolated (in flops) between the sampled size classes, clamped outside."""
    pts = sorted((works[f], works[f] / t) for f, t in samples.items())
    ws, rs = [a for a, _ in pts], [b for _, b in pts]
    return float(sum(w / np.interp(w, ws, rs) for w in works))


def run_reference(args, rank: int, world: int) -> None:
    if rank != 0:
        return
    cores = host_cores()
    omp = OpenMP()
    p = make_workload(args.workload, use_gpu_matmul=False)
    ref, kind, T_ref, works = reference_setup(p)  # libgomp reads OMP_NUM_THREADS (set at the top of this file) here
    omp.set(cores)
    omp_threads = omp.get()
    small, large = int(np.argmin(works)), int(np.argmax(works))
    frags = sorted({small, large}, key=lambda f: works[f])  # one sample per size class: smallest and largest fragment
    if args.warmup > 0:
        reference_sample_step(ref, kind, p, T_ref, small)
    n_samples = max(1, min(args.steps, 2))  # bounded: <= 3 timed samples whatever --steps says
    times = {f: [] for f in frags}
    for _ in range(n_samples):
        for f in frags:
            times[f].append(reference_sample_step(ref, kind, p, T_ref, f)[0])
    best = {f: float(np.min(v)) for f, v in times.items()}
    serial = extrapolate_all_fragments(works, best)
    # the reference's own concurrency knob (ThreadPoolExecutor(n_threads) around the per-fragment worker,
    # eri_sparse_DF.py:649-653, GIL released in the extension): k fragments at once, cores/k OpenMP threads each
    pool = {}
    for k in (2, 4):
        if cores < 2 * k or ref is None:
            continue
        per = max(1, cores // k)

        def work(_i):
            omp.set(per)  # nthreads-var is per calling thread
            return reference_sample_step(ref, kind, p, T_ref, small)[0]

        from concurrent.futures import ThreadPoolExecutor

        t0 = time.perf_counter()
        with ThreadPoolExecutor(max_workers=k) as ex:
            list(ex.map(work, range(k)))
        dt = time.perf_counter() - t0
        pool[k] = {"omp_threads_each": per, "seconds_for_k_fragments": dt, "speedup_vs_serial": k * best[small] / dt}
    omp.set(cores)
    best_k = max(pool, key=lambda k: pool[k]["speedup_vs_serial"]) if pool else 1
    speedup = max(1.0, pool[best_k]["speedup_vs_serial"]) if pool else 1.0
    if speedup == 1.0:
        best_k = 1
    value = serial / speedup
    measured = {f"fragment_{f}": {"nmo": int(p.TAs[f].shape[1]), "seconds": best[f], "all_samples": times[f],
                                  "executed_gflops": works[f] / best[f] / 1e9} for f in frags}
    sample = (f"{n_samples} timed pass(es) over fragments {frags} (nmo {[int(p.TAs[f].shape[1]) for f in frags]}: "
              f"{', '.join(f'{best[f]:.2f} s' for f in frags)}), every one of the {len(p.TAs)} fragments charged its executed "
              f"flops at the rate of its size class -> {serial:.0f} s serial; best of the reference's thread-pool "
              f"configurations: k={best_k} x {cores // best_k} OpenMP threads (x{speedup:.2f}); "
              f"{'oracle/_ref = unmodified reference extension' if kind == 'reference' else 'oracle C port'}, "
              f"OpenMP reports {omp_threads} threads on {cores} cores")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": n_samples,
        "steps_requested": args.steps, "warmup": min(args.warmup, 1), "ms_per_step": value * 1e3,
        "higher_is_better": False, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": config_dict(p, args.gpus),
        "extrapolated": True,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "omp_threads": omp_threads, "kind": kind,
                         "sample": sample, "measured": measured, "serial_all_fragments_s": serial,
                         "thread_pool": pool, "best_config": {"n_threads": best_k, "omp_each": cores // best_k}},
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


def host_write_bandwidth(cores: int) -> dict | None:
    """Host DRAM write / copy bandwidth with ALL host cores (a fresh process: inside a rank the library's pool only
    has cores / local ranks threads), taken on rank 0 while the other ranks wait."""
    code = ("import json; from quemb_b200 import eri_sparse_DF as q; w, c = q.host_bandwidth(2 << 30); "
            "from quemb_b200 import _lib; import ctypes; n = ctypes.c_int(); _lib.lib().qemb_host_threads(ctypes.byref(n)); "
            "print(json.dumps({'write_gbs': w, 'copy_gbs': c, 'threads': n.value}))")
    env = dict(os.environ, QEMB_HOST_THREADS=str(cores))
    env.pop("LOCAL_WORLD_SIZE", None)
    try:
        out = subprocess.run([sys.executable, "-c", code], env=env, cwd=str(ROOT), capture_output=True, text=True, timeout=120)
        return json.loads(out.stdout.strip().splitlines()[-1])
    except Exception:  # noqa: BLE001
        return None


def run_b200(args, rank: int, world: int, local_rank: int) -> None:
    import torch

    from quemb_b200 import _lib, consumer
    from quemb_b200 import eri_sparse_DF as q
    from quemb_b200.driver import DistributedQueue, FragmentERIBuilder, fragment_cost, partition_lpt

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

    def reduce_ranks(x: float, op: str) -> float:
        if dist is None:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX if op == "max" else dist.ReduceOp.SUM)
        return float(t.item())

    def max_over_ranks(x: float) -> float:
        return reduce_ranks(x, "max")

    def sum_over_ranks(x: float) -> float:
        return reduce_ranks(x, "sum")

    peak_burst, peak_sust = fp64_peak_tflops(torch) if rank == 0 else (0.0, 0.0)
    cores = host_cores()
    try:  # torchrun exports OMP_NUM_THREADS=1: give numpy's BLAS (synthetic-data generation only) this rank's share
        from threadpoolctl import threadpool_limits

        threadpool_limits(limits=max(1, cores // world))
    except Exception:  # noqa: BLE001
        pass
    host_bw = host_write_bandwidth(cores) if rank == 0 else None
    barrier()

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
    # the consumer kernel right behind the path, informational: J/K of one density from the
    # (ij|kl) the last fragment left on the device (HBM bound: 8 npair^2 bytes read once)
    consumer_kernel = None
    if rank == 0 and mine:
        n_last = p.TAs[mine[-1]].shape[1]
        Dm = np.random.default_rng(1).standard_normal((n_last, n_last))
        Dm = Dm + Dm.T
        best, nb = None, 0.0
        for _ in range(4):
            b.ws.jk(Dm)
            ms_jk, nb = b.ws.jk_time()
            best = ms_jk if best is None else min(best, ms_jk)
        consumer_kernel = {"kernel": "jk_stream_kernel + jk_reduce_kernel (J and K of one density from the resident (ij|kl))",
                           "fragment_nmo": int(n_last), "ms": best, "gbs": nb / best / 1e6}

    # ---- e2e: host buffers in, pinned host buffers out ------------------------------------------------

    # How a host-bound (ij|kl) travels: the lower triangle always crosses PCIe; of the upper triangle a fraction crosses
    # PCIe as well and the rest is rebuilt by the library's host threads.  "auto": every rank measures its own copy
    # rate and mirror rate at the same moment (all ranks, mutual contention included) and balances the two
    # (qemb_ctx_autotune_output); "triangle" / "full" force the two extremes.
    out_mode = args.e2e_output
    barrier()
    if out_mode == "auto":
        try:
            tuned = b.tune_output()
        except Exception as e:  # noqa: BLE001  -- tuning must never take the measurement down: fall back to a fixed mode
            ctx.set_output_mode(world == 1)
            ctx.set_full_column_fraction(0.0)
            tuned = {"error": f"{type(e).__name__}: {e}", "upper_fraction_over_pcie": 0.0 if world == 1 else 1.0}
    else:
        ctx.set_output_mode(out_mode == "triangle")
        ctx.set_full_column_fraction(0.0)
        tuned = {"d2h_gbs": None, "mirror_gbs": None, "upper_fraction_over_pcie": 0.0 if out_mode == "triangle" else 1.0}
    tuned_all = [tuned]
    if dist is not None:
        tuned_all = [None] * world
        dist.all_gather_object(tuned_all, tuned)
    checks = []
    order_all = sorted(range(nfrag), key=lambda i: (-costs[i], i))
    taken_counts = []
    qcount = [0]

    def sink(i, G):
        checks.append(float(G[0, 0]) + float(G[-1, -1]))  # the step's result is read on the host

    tl = {"upload+stage_s": 0.0, "build_s": 0.0}

    def stage_from_host():
        # everything a caller holding HOST arrays pays before the first fragment: (P|mu nu), L, S_abs up, metric applied
        if dist is None:
            work_t.upload(p.unique_dense_data)
            b.stage(None, None, None, p.L_PQ, p.S_abs, whiten=True, tensor=work_t)
        else:  # one host->device pass of the tensor per NODE: 1/N of the columns per rank, exchanged over NVLink
            b.stage_distributed(p.unique_dense_data, p.L_PQ, p.S_abs, dist, tensor=work_t)

    def e2e_step():
        ta = time.perf_counter()
        stage_from_host()
        tb = time.perf_counter()
        if dist is None:
            b.build(p.TAs, p.MO_coeff_epsilon, sink, indices=mine)
        else:
            # fragments are handed out dynamically across the ranks (largest first, an atomic counter in the process
            # group's store): the ranks of one box do not drain their outputs at the same rate (PCIe topology, host
            # share), and delivery, not the kernels, bounds this number
            qcount[0] += 1
            taken = b.build_dynamic(p.TAs, p.MO_coeff_epsilon, sink, DistributedQueue(dist, order_all, f"e2e{qcount[0]}"))
            taken_counts.append(len(taken))
        tl["upload+stage_s"] += tb - ta
        tl["build_s"] += time.perf_counter() - tb

    e2e_steps = max(1, args.steps if args.e2e_steps <= 0 else args.e2e_steps)
    e2e_step()
    ctx.synchronize(); barrier()
    tl = {k: 0.0 for k in tl}
    xfer0 = ctx.transfer_bytes
    ctx.mark(2, "compute")
    tw = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    ctx.mark(3, "copy")
    ctx.synchronize(); barrier()
    wall_e2e = (time.perf_counter() - tw) / e2e_steps
    xfer1 = ctx.transfer_bytes
    # bytes the copy engines actually moved per step (counted by the library)
    h2d, d2h_moved = (xfer1[0] - xfer0[0]) / e2e_steps, (xfer1[1] - xfer0[1]) / e2e_steps
    t_e2e = max_over_ranks(ctx.mark_elapsed_ms(2, 3) * 1e-3) / e2e_steps
    wall_e2e = max_over_ranks(wall_e2e)
    tl_max = {k: max_over_ranks(v / e2e_steps) for k, v in tl.items()}
    frag_taken = [len(mine)]
    if dist is not None:
        frag_taken = [None] * world
        dist.all_gather_object(frag_taken, taken_counts[-1] if taken_counts else 0)
    clocks = sampler.stop() if sampler else None
    h2d_all, d2h_all = sum_over_ranks(float(h2d)), sum_over_ranks(float(d2h_moved))
    result_all = float(sum(8 * (t.shape[1] * (t.shape[1] + 1) // 2) ** 2 for t in p.TAs))
    launches_all = int(sum_over_ranks(float(launches)))

    # aggregate device->host rate the box can absorb (all ranks copy 2 GiB to pinned memory at once): the delivery
    # roofline of e2e next to the host-DRAM write bandwidth
    nprobe = 1 << 28
    src = torch.empty(nprobe, dtype=torch.float64, device="cuda")
    dst = torch.empty(nprobe, dtype=torch.float64, pin_memory=True)
    dst.copy_(src); barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); dst.copy_(src, non_blocking=True); e1.record(); torch.cuda.synchronize()
    d2h_gbs_all = sum_over_ranks(nprobe * 8 / (max_over_ranks(e0.elapsed_time(e1)) * 1e-3) / 1e9)
    del src, dst

    # ---- e2e_consumer: same inputs from the host, consumers evaluated on the resident (ij|kl) ----------
    rng = np.random.default_rng(11)
    hf_dm = rng.standard_normal((p.nao, p.nao)); hf_dm = hf_dm + hf_dm.T
    hf_veff = rng.standard_normal((p.nao, p.nao)); hf_veff = hf_veff + hf_veff.T
    S_ao = np.eye(p.nao)
    rdms = {}
    for i in mine:
        n = p.TAs[i].shape[1]
        c = np.linalg.qr(rng.standard_normal((n, max(1, n // 3))))[0]
        rdms[i] = c @ c.T
    n_frags = {i: max(1, p.TAs[i].shape[1] // 2) for i in mine}
    cons_out = {}

    def consumer_step():
        stage_from_host()
        cons_out.update(consumer.initialize_fragments(b, p.TAs, p.MO_coeff_epsilon, hf_dm, S_ao, hf_veff, rdms, n_frags,
                                                      indices=mine))

    consumer_step()
    ctx.synchronize(); barrier()
    xc0 = ctx.transfer_bytes
    cons_steps = max(1, min(e2e_steps, 5))
    ctx.mark(4, "compute")
    for _ in range(cons_steps):
        consumer_step()
    ctx.mark(5, "compute")
    ctx.synchronize(); barrier()
    xc1 = ctx.transfer_bytes
    t_cons = max_over_ranks(ctx.mark_elapsed_ms(4, 5) * 1e-3) / cons_steps
    cons_h2d = sum_over_ranks((xc1[0] - xc0[0]) / cons_steps)
    cons_d2h = sum_over_ranks((xc1[1] - xc0[1]) / cons_steps)
    cons_check = float(sum(float(np.sum(v["e2"])) for v in cons_out.values()))

    # ---- e2e_file: bounded sample of delivery-to-file (background writer fed from the pinned ring) -------
    e2e_file = None
    if rank == 0 and not args.no_file:
        from quemb_b200.eri_file import AsyncDatasetWriter, NpyDirFile

        sample = mine[: args.file_fragments]
        nbytes = sum(8 * (p.TAs[i].shape[1] * (p.TAs[i].shape[1] + 1) // 2) ** 2 for i in sample)
        base = Path(args.file_dir or tempfile.gettempdir())
        free = shutil.disk_usage(base).free
        if free < 2 * nbytes + (4 << 30):
            e2e_file = {"skipped": f"{free / 1e9:.0f} GB free under {base}, sample needs {nbytes / 1e9:.1f} GB"}
        else:
            d = Path(tempfile.mkdtemp(prefix="qemb_eri_", dir=base))
            try:
                fh = NpyDirFile(d / "eri_file.h5", "w")
                tf = time.perf_counter()
                with AsyncDatasetWriter(fh) as w:
                    b.build(p.TAs, p.MO_coeff_epsilon, lambda i, G: w.submit(f"f{i}", G), indices=sample)
                dt = time.perf_counter() - tf
                back = np.load(d / "eri_file.h5.d" / f"f{sample[0]}.npy", mmap_mode="r")
                ok = bool(back.shape[0] == p.TAs[sample[0]].shape[1] * (p.TAs[sample[0]].shape[1] + 1) // 2)
                e2e_file = {"seconds": dt, "bytes": nbytes, "gbs": nbytes / dt / 1e9, "fragments": len(sample),
                            "writer_busy_s": w.busy_s, "container": "NpyDirFile (h5py absent on this box)",
                            "readback_shape_ok": ok, "dir": str(base),
                            "note": "bounded sample: the first fragments of rank 0 through build() -> AsyncDatasetWriter -> "
                                    "create_dataset(f{i}); page cache included (no fsync)"}
            finally:
                shutil.rmtree(d, ignore_errors=True)
    barrier()

    # ---- cpu baseline (rank 0, N = 1 only): bounded sample of the reference's CPU path ------------------
    cpu_baseline = None
    int8_extra = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        ref, kind, T_ref, works = reference_setup(p)
        small, large = int(np.argmin(works)), int(np.argmax(works))
        frags = sorted({small, large}, key=lambda f: works[f])
        best, parity = {}, {}
        for f in frags:
            dt, G_ref = reference_sample_step(ref, kind, p, T_ref, f)
            best[f] = dt
            # the same fragment through the B200 path (whitened tensor as timed above): full-size parity, untimed
            got = {}
            b.build(p.TAs, p.MO_coeff_epsilon, lambda i, G: got.__setitem__(i, float(np.abs(G - G_ref).max())), indices=[f])
            parity[f] = got.get(f)
            if f == large and not args.no_int8_extra:
                # EXTRA, never the value (north star: all FP64): the opt-in int8-slice K3 in its guaranteed-tolerance
                # mode (smallest slice count whose worst-case bound meets 1e-10 absolute on every (ij|kl), FP64 kernel
                # otherwise), checked on the same full matrix of the largest fragment against the compiled reference
                try:
                    ctx.set_k3_int8_tolerance(args.int8_tolerance)
                    got8 = {}
                    b.build(p.TAs, p.MO_coeff_epsilon, lambda i, G: got8.__setitem__(i, float(np.abs(G - G_ref).max())), indices=[f])
                    choice = b.ws.k3_choice
                    ctx.synchronize()
                    ctx.mark(6, "compute")
                    for i in mine:
                        b.build_device(p.TAs, p.MO_coeff_epsilon, Gdev.data_ptr(), indices=[i])
                    ctx.mark(7, "compute")
                    ctx.synchronize()
                    int8_extra = {"mode": f"guaranteed tolerance {args.int8_tolerance:g} absolute (qemb_ctx_set_k3_int8_tolerance)",
                                  "slices_chosen_largest_fragment": choice[0], "max_column_exponent": choice[1],
                                  "parity_max_abs_vs_reference_largest_fragment": got8.get(f),
                                  "all_fragments_device_s": ctx.mark_elapsed_ms(6, 7) * 1e-3,
                                  "note": "metric application not included (the tensor is already whitened); compare with "
                                          "`value` minus ~0.03 s"}
                finally:
                    ctx.set_k3_int8_tolerance(0.0)
            del G_ref
        serial = extrapolate_all_fragments(works, best)
        cpu_baseline = {
            "value": serial, "unit": UNIT, "cores": cores, "kind": kind, "extrapolated": True,
            "sample": (f"fragments {frags} (nmo {[int(p.TAs[f].shape[1]) for f in frags]}): "
                       f"{', '.join(f'{best[f]:.2f} s' for f in frags)} (one pass each), every one of the {nfrag} fragments "
                       f"charged its executed flops at the rate of its size class; fragments serial, OpenMP over {cores} "
                       "threads (the thread-pool configurations are timed by --impl reference)"),
            "parity_max_abs": max(v for v in parity.values() if v is not None),
            "parity_per_fragment": {f"fragment_{f}_nmo_{int(p.TAs[f].shape[1])}": parity[f] for f in frags},
            "parity_note": "max |(ij|kl)_B200 - (ij|kl)_reference| over the FULL npair^2 matrix of the smallest and the "
                           "largest fragment of the workload; bar 1e-10",
        }

    # ---- sharded extra (N > 1): one oversize fragment, auxiliary index split over the ranks ---------------
    sharded = None
    if dist is not None and not args.no_sharded_extra:
        del Gdev, raw
        b.close()
        work_t = None
        torch.cuda.empty_cache()
        try:
            sharded = sharded_extra(args, torch, dist, q, rank, world, local_rank, peak_sust)
        except Exception as e:  # noqa: BLE001  -- an extra must never take the headline line down
            sharded = {"error": f"{type(e).__name__}: {e}"}

    if rank != 0:
        if dist is not None:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (K3, FP64 tensor pipe) ------------------------------------------
    n_k3 = max(klaunch["K3"], 1)
    k3_ms = ktimes["K3"] / n_k3
    k3_flops = kwork.get("F_K3", 0.0) / n_k3
    achieved = k3_flops / (k3_ms * 1e-3) / 1e12 if k3_ms > 0 else 0.0
    traffic, traffic_src = None, None
    tj = ROOT / "profiles" / "k3_traffic_r2.json"  # ncu pass over THIS workload's K3 launches: sum / launches
    if tj.exists() and args.workload == "c32h66" and world == 1 and not args.k3_int8:
        try:
            rec = json.loads(tj.read_text())
            traffic, traffic_src = rec.get("dram_bytes_per_launch_mean"), rec.get("source")
        except Exception:  # noqa: BLE001
            traffic = None
    per_kernel = {}
    for k, fk, bk in (("K1", "F_K1", "B_K1"), ("K2a", "F_K2a", "B_K2a"), ("K3", "F_K3", "B_K3")):
        if ktimes[k] > 0:
            per_kernel[k] = {"ms_total": ktimes[k], "tflops": kwork.get(fk, 0) / (ktimes[k] * 1e-3) / 1e12,
                             "gbs": kwork.get(bk, 0) / (ktimes[k] * 1e-3) / 1e9,
                             "frac_fp64_peak": kwork.get(fk, 0) / (ktimes[k] * 1e-3) / 1e12 / peak_sust if peak_sust else None}
    peaks = {}
    mp = ROOT / "MEASURED_PEAKS.json"
    if mp.exists():
        peaks = json.loads(mp.read_text())
    if consumer_kernel and peaks.get("hbm_gbs"):
        consumer_kernel["hbm_frac"] = consumer_kernel["gbs"] / peaks["hbm_gbs"]
    host_write = (host_bw or {}).get("write_gbs")
    line = {
        "metric": METRIC, "value": t_dev, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": t_dev * 1e3, "higher_is_better": False, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": config_dict(p, world),
        "detail": {"fragments_rank0": len(mine),
                   "metric_application": ("split over the ranks by columns, exchanged with NCCL broadcasts over NVLink"
                                          if world > 1 else "one rank"),
                   "k3": (f"opt-in int8-slice K3, {args.k3_int8} slices (roofline.achieved counts the FP64-equivalent flops)"
                          if args.k3_int8 else "FP64 DMMA (default)")},
        "fragments_per_s": nfrag / t_dev,
        "roofline": {"bound": "tensor", "kernel": "gram_kernel (K3: G = X^T X, FP64 DMMA.8x8x4)",
                     "achieved": achieved, "peak": peak_sust, "unit": "TFLOP/s",
                     "frac": achieved / peak_sust if peak_sust else None, "traffic": traffic, "traffic_source": traffic_src,
                     "peak_source": "cuBLAS FP64 DGEMM 8192^3 measured in this run, sustained "
                                    f"(burst {peak_burst:.1f}); MEASURED_PEAKS.json has no FP64 entry",
                     "algorithmic_flops_per_launch": k3_flops, "avg_launch_ms": k3_ms, "launches_measured": n_k3},
        "per_kernel": per_kernel,
        "hbm_peak_gbs_measured": peaks.get("hbm_gbs"),
        "staging_s": staging_s,
        "consumer_kernel": consumer_kernel,
        "cpu_baseline": cpu_baseline,
        "e2e": {"value": t_e2e, "unit": UNIT, "h2d_bytes_per_step": int(h2d_all), "d2h_bytes_per_step": int(d2h_all),
                "result_bytes_per_step": int(result_all), "wall_s": wall_e2e, "steps": e2e_steps,
                "output_mode": out_mode,
                "autotune_per_rank": tuned_all,
                "fragments_taken_per_rank_last_step": frag_taken,
                "timeline_max_over_ranks": tl_max,
                "host": {"cores": cores, "threads_per_rank": q_host_threads(_lib), "dram_write_gbs": host_write,
                         "dram_copy_gbs": (host_bw or {}).get("copy_gbs"), "d2h_aggregate_gbs": d2h_gbs_all},
                # every (ij|kl) matrix must be WRITTEN into host DRAM once: result bytes / write bandwidth / e2e
                "host_roofline_frac": (result_all / 1e9 / host_write / t_e2e) if host_write else None,
                # the PCIe side of the same delivery: bytes that crossed / aggregate D2H rate / e2e
                "d2h_roofline_frac": (d2h_all / 1e9 / d2h_gbs_all / t_e2e) if d2h_gbs_all else None,
                "kernel_fraction": t_dev / t_e2e if t_e2e else None,
                "note": ("includes upload of (P|mu nu)+L+S_abs from pageable host memory"
                         + (" (1/N of the columns per rank, shares exchanged over NVLink)" if world > 1 else "")
                         + ", metric application, all fragments"
                         + (" (handed out dynamically across the ranks)" if world > 1 else "")
                         + ", every full npair x npair G delivered in pinned host memory (lower triangle over PCIe; of the "
                         "upper triangle the per-rank fraction autotune_per_rank[r].upper_fraction_over_pcie crosses PCIe "
                         "too, the rest is rebuilt by the library's host threads) inside the timed region")},
        "e2e_consumer": {"value": t_cons, "unit": UNIT, "h2d_bytes_per_step": int(cons_h2d), "d2h_bytes_per_step": int(cons_d2h),
                         "steps": cons_steps, "checksum": cons_check,
                         "note": "same host inputs, same transforms; get_veff + update_ebe_hf two-electron sums of every "
                                 "fragment evaluated on the resident (ij|kl) (FragmentERIBuilder.build_jk): only nmo x nmo "
                                 "matrices return to the host"},
        "e2e_file": e2e_file,
        "sharded": sharded,
        "k3_int8_extra": int8_extra,
        "gpu_launches": launches_all,
        "clocks": clocks,
    }
    print(json.dumps(line), flush=True)
    if dist is not None:
        dist.destroy_process_group()


def q_host_threads(_lib) -> int:
    n = ctypes.c_int()
    _lib.lib().qemb_host_threads(ctypes.byref(n))
    return n.value


def sharded_extra(args, torch, dist, q, rank: int, world: int, local_rank: int, peak_tf: float) -> dict | None:
    """BASELINE configs[3] / north star "only fragments whose B tensor exceeds one GPU shard the auxiliary index P and
    combine with an NCCL reduce-sum over NVLink": (a) parity of the sharded build against the ORACLE on a companion
    case the CPU finishes in seconds, (b) time of oversize fragments (nmo 400 / naux 8000: G = 51 GB each)."""
    from quemb_b200 import synthetic as syn
    from quemb_b200.sharded import AuxShardedBuilder

    class DeviceSink(AuxShardedBuilder):
        def _to_host(self, G, npair):  # keep the summed matrix on the owner's device
            return G

    def sync():
        torch.cuda.synchronize()
        dist.barrier()
        torch.cuda.synchronize()

    out = {"variant": "AuxShardedBuilder: slab of L^-1 (P|mu nu) per rank, partial (ij|kl) per rank, asynchronous NCCL "
                      "reduce-sum onto a rotating owner (deterministic); handoff to NCCL by CUDA event, no host sync",
           "world": world}
    # (a) companion parity case against the oracle
    pc = syn.sweep_point(100, 2000, max_bytes=1e9, matmul=gpu_matmul)
    bb = AuxShardedBuilder(dist, device=local_rank)
    bb.stage(pc.unique_dense_data, (pc.naux, pc.nao, pc.nao), pc.exch_reachable, pc.L_PQ, pc.S_abs)
    frs = [len(pc.TAs) - 1, 0][: max(1, min(2, world))]
    errs = {}

    def sink(i, G):
        from oracle import oracle

        ref = oracle.reference_module()
        OpenMP().set(max(1, host_cores() // world))  # torchrun exports OMP_NUM_THREADS=1
        if ref is not None:
            T = ref.SemiSparseSym3DTensor(pc.unique_dense_data, (pc.naux, pc.nao, pc.nao), [x.tolist() for x in pc.exch_reachable])
            want = ref.transform_integral(T, pc.TAs[i], pc.S_abs, pc.L_PQ, pc.MO_coeff_epsilon)
        else:
            want = oracle.transform_integral(pc.unique_dense_data, pc.exch_reachable, pc.TAs[i], pc.S_abs, pc.L_PQ,
                                             pc.MO_coeff_epsilon)
        errs[i] = float(np.abs(G - want).max())

    bb.build(pc.TAs, pc.MO_coeff_epsilon, sink, indices=frs)
    sync()
    gathered = [None] * world
    dist.all_gather_object(gathered, errs)
    allerr = {k: v for g in gathered for k, v in g.items()}
    out["parity_vs_oracle"] = {"case": f"{pc.name}: nao {pc.nao}, naux {pc.naux}, nmo {pc.TAs[-1].shape[1]}, slabs of "
                                       f"{-(-pc.naux // world)} auxiliary rows", "fragments": len(allerr),
                               "max_abs": max(allerr.values()) if allerr else None, "bar": 1e-10}
    del bb
    torch.cuda.empty_cache()
    # (b) oversize fragments
    t0 = time.perf_counter()
    p = syn.sweep_point(args.sharded_nmo, args.sharded_naux, max_bytes=args.sharded_max_gb * 1e9, matmul=gpu_matmul,
                        device_linalg=True)
    gen_s = time.perf_counter() - t0
    nf = args.sharded_fragments
    frags = [len(p.TAs) - 1 - k for k in range(nf)]
    nmo = p.TAs[frags[0]].shape[1]
    npair = nmo * (nmo + 1) // 2
    b = DeviceSink(dist, device=local_rank)
    t0 = time.perf_counter()
    b.stage(p.unique_dense_data, (p.naux, p.nao, p.nao), p.exch_reachable, p.L_PQ, p.S_abs)
    sync()
    stage_s = time.perf_counter() - t0
    b.build(p.TAs, p.MO_coeff_epsilon, None, indices=frags)  # warm-up (buffers, plans)
    sync()
    t0 = time.perf_counter()
    b.build(p.TAs, p.MO_coeff_epsilon, None, indices=frags)
    sync()
    t = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    secs = float(t.item())
    flops = float(p.naux) * npair * (npair + 1) * nf
    out["oversize"] = {"case": f"{p.name}: nao {p.nao}, naux {p.naux}, nmo {nmo}, npair {npair}, (ij|kl) "
                               f"{8 * npair * npair / 1e9:.1f} GB per fragment", "fragments": nf, "seconds": secs,
                       "seconds_per_fragment": secs / nf, "K3_tflops_aggregate": flops / secs / 1e12,
                       "frac_of_N_x_fp64_peak": (flops / secs / 1e12) / (world * peak_tf) if peak_tf else None,
                       "reduce_bytes_per_fragment": 8 * npair * npair, "slab_rows": b.slab[1] - b.slab[0],
                       "slab_tensor_gb": b.tensor.nonzero_size * 8 / 1e9, "full_tensor_gb": p.unique_dense_data.nbytes / 1e9,
                       "stage_s": stage_s, "gen_s": gen_s,
                       "timing": "wall clock between barrier+cudaDeviceSynchronize pairs, max over ranks; summed G stays on "
                                 "the owners' devices"}
    return out


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="c32h66", choices=["c32h66", "octane", "tiny"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--e2e-output", default="auto", choices=["auto", "triangle", "full"],
                    help="how a host-bound (ij|kl) travels: lower triangle + host mirror, the full square, or (auto, default) "
                         "whatever the library's self-tuning measures to be fastest on this host, per rank")
    ap.add_argument("--e2e-steps", type=int, default=0, help="timed e2e steps (default: --steps)")
    ap.add_argument("--no-file", action="store_true", help="skip the bounded delivery-to-file sample")
    ap.add_argument("--file-dir", default=None)
    ap.add_argument("--file-fragments", type=int, default=2)
    ap.add_argument("--no-sharded-extra", action="store_true")
    ap.add_argument("--no-int8-extra", action="store_true", help="skip the opt-in int8 K3 extra (N = 1, with the cpu baseline)")
    ap.add_argument("--int8-tolerance", type=float, default=1e-10)
    ap.add_argument("--sharded-nmo", type=int, default=400)
    ap.add_argument("--sharded-naux", type=int, default=8000)
    ap.add_argument("--sharded-max-gb", type=float, default=6.0)
    ap.add_argument("--sharded-fragments", type=int, default=2)
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
