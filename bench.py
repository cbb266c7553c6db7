#!/usr/bin/env python
"""Headline benchmark: interior-point iterations per second on the synthetic block
multi-cone model of BASELINE.json config 5 (Schur dimension N, q = 2N: Nonnegative /
EpiNormEucl(8) / EpiPerSquare(8) / PosSemidefTri(side 16) quarters), with the
Schur-assemble + Cholesky milliseconds and the FP64-tensor roofline of the dominant
kernel (G'HG syrk on DMMA) beside it.

  python bench.py --gpus 1 --steps 5 --warmup 3            # this repo's CUDA path
  python bench.py --impl reference --steps 3 --warmup 1    # CPU restatement of the reference on host cores

One "step" = one interior-point iteration (update_lhs -> 4 directions with refinement ->
line search), the unit the reference itself counts (Solvers.jl:371-373, num_iters).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def _host_cores():
    """(physical cores, logical cpus) of this box."""
    logical = os.cpu_count() or 1
    try:
        import psutil

        return int(psutil.cpu_count(logical=False) or logical), logical
    except Exception:
        return logical, logical


# torchrun exports OMP_NUM_THREADS=1 to every rank; OpenBLAS reads it when numpy is imported, which
# would time the CPU arms on ONE thread.  Rank 0 (the only rank that does host LAPACK / runs the CPU
# arm) gets the physical cores, stated in the JSON line; the protocol of the reference's own
# benchmarks states its BLAS thread count the same way (benchmarks/natvsext/run.jl:38-60).
PHYS_CORES, LOGICAL_CPUS = _host_cores()
if int(os.environ.get("RANK", "0")) == 0:
    for _v in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS"):
        os.environ[_v] = str(PHYS_CORES)

import numpy as np  # noqa: E402


def blas_threads():
    """BLAS threads actually in use (threadpoolctl), as reported in cpu_baseline.cores."""
    try:
        from threadpoolctl import threadpool_info

        n = [int(i.get("num_threads", 0)) for i in threadpool_info() if i.get("user_api") == "blas"]
        return max(n) if n else 1
    except Exception:
        return 1


def set_blas_threads(n):
    try:
        from threadpoolctl import threadpool_limits

        threadpool_limits(int(n), user_api="blas")
    except Exception:
        pass


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--N", type=int, default=int(os.environ.get("HYB_BENCH_N", 8192)), help="Schur dimension of the cfg5 model")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""

    FIELDS = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index=0):
        self.index = index
        self.samples = []
        self.source = "nvidia-smi --query-gpu"
        self._stop = threading.Event()
        self._thread = None

    def _run_nvml(self) -> bool:
        """In-process NVML queries (the library nvidia-smi itself reads): no process spawn and no enumeration of
        every GPU of the node five times a second beside the timed iterations.  False if NVML is unavailable."""
        try:
            import pynvml as nv

            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            smax = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            bits = {"hw_slowdown": nv.nvmlClocksEventReasonHwSlowdown if hasattr(nv, "nvmlClocksEventReasonHwSlowdown") else nv.nvmlClocksThrottleReasonHwSlowdown,
                    "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", None) or nv.nvmlClocksThrottleReasonHwThermalSlowdown,
                    "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", None) or nv.nvmlClocksThrottleReasonSwThermalSlowdown,
                    "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", None) or nv.nvmlClocksThrottleReasonSwPowerCap}
            get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
            nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
            get_reasons(h)
        except Exception:
            return False
        self.source = "NVML in-process (nvmlDeviceGetClockInfo / nvmlDeviceGetCurrentClocksEventReasons, what nvidia-smi reads)"
        while not self._stop.is_set():
            try:
                sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
                r = get_reasons(h)
                try:
                    pw = nv.nvmlDeviceGetPowerUsage(h) / 1000.0
                except Exception:
                    pw = 0.0
                self.samples.append([str(sm), str(smax), str(pw)] + ["Active" if (r & bits[k]) else "Not Active" for k in
                                                                       ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")])
            except Exception:
                pass
            self._stop.wait(0.1)
        return True

    def _run(self):
        if os.environ.get("HYB_CLOCKS_NVIDIA_SMI") != "1" and self._run_nvml():
            return
        while not self._stop.is_set():
            try:
                out = subprocess.run(
                    ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits"],
                    capture_output=True, text=True, timeout=5,
                ).stdout.strip()
                if out:
                    self.samples.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            self._stop.wait(0.2)

    def start(self):
        self._thread = threading.Thread(target=self._run, daemon=True)
        self._thread.start()

    def stop(self):
        self._stop.set()
        if self._thread:
            self._thread.join(timeout=6)
        sm, smax, reasons = [], [], set()
        for s in self.samples:
            try:
                sm.append(float(s[0]))
                smax.append(float(s[1]))
            except Exception:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), s[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "source": self.source}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(smax)), "reasons": sorted(reasons), "samples": len(sm),
                "source": self.source}


# Solver options of the reference's own benchmark protocol (benchmarks/natvsext/run.jl:44-60), used by both
# arms.  At the default tolerances (eps^0.5 = 1.5e-8) this model stalls a hair short of the feasibility
# tolerance in the CPU restatement and on the GPU alike (SlowProgress -> NearOptimal after 23 iterations,
# tests/golden/trajectories_big.json); under the benchmark tolerances both end Optimal after 15.
BENCH_TOL = dict(tol_rel_opt=1e-7, tol_feas=1e-7, tol_abs_opt=1e-10, tol_infeas=1e-10)


def build_model(N):
    import hypatia_b200  # noqa: F401
    from hypatia_b200 import instances

    return instances.block_multicone(N=N, seed=1)


def workload_config(N, model, extra=None):
    cfg = {
        "workload": f"cfg5 synthetic block multi-cone model: Schur dim N={N}, q={model.q}, "
                    f"{len(model.cones)} cones (Nonnegative / EpiNormEucl(8) / EpiPerSquare(8) / PosSemidefTri(side 16) quarters), Float64",
        "step": "one interior-point iteration (CombinedStepper, QRChol system solver, reference defaults)",
        "solver_options": "tolerances of the reference's benchmark protocol (benchmarks/natvsext/run.jl:44-60): "
                          "tol_rel_opt = tol_feas = 1e-7, tol_abs_opt = tol_infeas = 1e-10; everything else default",
        "N": N, "q": int(model.q), "cones": len(model.cones),
        "l2": "inputs larger than L2 (G*Q2 is %.2f GB, Schur matrix %.2f GB)" % (model.q * N * 8 / 1e9, N * N * 8 / 1e9),
    }
    if extra:
        cfg.update(extra)
    return cfg


class IterTimer:
    """Per-iteration wall clock + library statistics, hooked on solver.iter_callback."""

    def __init__(self, ctx=None, sync=None):
        self.ctx, self.sync = ctx, sync
        self.rows = []
        self._t = None
        self._ph = {}

    def start(self):
        self._ph = {}  # the driver zeroes its phase clocks at the start of every solve
        if self.sync:
            self.sync()
        if self.ctx:
            self.ctx.stats(reset=True)
            self._launch0 = self.ctx.launch_count()
        self._t = time.perf_counter()

    def __call__(self, solver):
        if self.sync:
            self.sync()
        now = time.perf_counter()
        row = {"wall_s": now - self._t, "n_cand": solver.stepper.last_num_cand, "n_ref": solver.stepper.last_num_ref}
        for name in ("upsys", "uprhs", "getdir", "search"):  # the driver's own phase clocks (Solvers.jl:529-564)
            tot = getattr(solver, "time_" + name, 0.0)
            row["ph_" + name] = 1e3 * (tot - self._ph.get(name, 0.0))
            self._ph[name] = tot
        if self.ctx:
            dev_ms, h2d, d2h = self.ctx.stats(reset=True)
            tw, ts, tf = self.ctx.last_schur_timings()
            if getattr(self, "shard_ctx", None):  # root/worker form: rank 0 assembles its share in a second context
                sc = self.shard_ctx()
                d2, a2, b2 = sc.stats(reset=True)
                tw, ts, _ = sc.last_schur_timings()
                dev_ms, h2d, d2h = dev_ms + d2, h2d + a2, d2h + b2
            if hasattr(solver.syssolver, "last_assemble_ms"):  # single-process multi-GPU: slowest device's assembly
                tw, ts = 0.0, solver.syssolver.last_assemble_ms
                dev_ms += ts  # the shards' kernels run in contexts of their own: count the slowest one's device time
            lc = self.ctx.launch_count()
            red = getattr(solver.syssolver, "time_reduce_ms", 0.0)
            row.update(dev_ms=dev_ms, h2d=h2d, d2h=d2h, t_w=tw, t_syrk=ts, t_fact=tf, launches=lc - self._launch0,
                       reduce_ms=red - getattr(self, "_red0", 0.0))
            self._red0 = red
            self._launch0 = lc
        self.rows.append(row)
        self._t = time.perf_counter()


def cublas_dgemm_tflops(n=8192, reps=3):
    """Measuring stick only (never on the hot path): FP64 GEMM rate of cuBLAS on this GPU."""
    import torch

    a = torch.randn(n, n, dtype=torch.float64, device="cuda")
    b = torch.randn(n, n, dtype=torch.float64, device="cuda")
    torch.matmul(a.T, b)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        torch.matmul(a.T, b)
    e1.record()
    torch.cuda.synchronize()
    del a, b
    torch.cuda.empty_cache()
    return 2.0 * n**3 * reps / (e0.elapsed_time(e1) * 1e-3) / 1e12


def ncu_traffic(N, K):
    """DRAM bytes per launch of the syrk kernel from the committed ncu capture (None if the
    shape was not captured)."""
    path = os.path.join(ROOT, "profiles", "r02_syrk_N8192_ncu_full.json")
    try:
        d = json.load(open(path))
        if f"N={N}, K={K}," in d["kernel"]:
            return d["dram_bytes_per_launch"]
    except Exception:
        pass
    return None


def run_cpu_port(model, steps, warmup, preset_fact=None, budget_s=150.0):
    """CPU restatement of the reference path (oracle/, numpy/scipy -> OpenBLAS) on host cores:
    `warmup` untimed + up to `steps` timed interior-point iterations, cut short (never below one
    timed iteration) once the iteration loop has used `budget_s` seconds."""
    from hypatia_b200.host.solver import Solver
    from oracle.cones import CpuConeSet
    from oracle.qrchol import QRCholDenseSystemSolver

    set_blas_threads(PHYS_CORES)
    ss = QRCholDenseSystemSolver()
    cpu = Solver(CpuConeSet, ss, iter_limit=warmup + steps, **BENCH_TOL)
    cpu.preset_AG_fact = preset_fact
    timer = IterTimer()
    loop_t0 = [None]

    def cb(solver):
        timer(solver)
        if loop_t0[0] is None:
            loop_t0[0] = time.perf_counter() - timer.rows[-1]["wall_s"]
        if time.perf_counter() - loop_t0[0] > budget_s and len(timer.rows) > warmup:
            solver.iter_limit = solver.num_iters  # bounded sample: stop after this iteration

    cpu.iter_callback = cb
    cpu.load(model)
    t0 = time.perf_counter()
    orig_init = cpu._initialize_point

    def init_and_start():
        orig_init()
        timer.start()

    cpu._initialize_point = init_and_start
    cpu.solve()
    rows = timer.rows[warmup:] if len(timer.rows) > warmup else timer.rows
    wall = sum(r["wall_s"] for r in rows)
    return {
        "iters_per_s": len(rows) / wall if wall > 0 else 0.0,
        "steps": len(rows),
        "ms_per_step": 1e3 * wall / max(len(rows), 1),
        "assemble_ms": 1e3 * ss.time_assemble / max(cpu.num_iters, 1),
        "factor_ms": 1e3 * cpu.time_upfact / max(cpu.num_iters, 1),
        "total_s": time.perf_counter() - t0,
        "status": cpu.status,
        "threads": blas_threads(),
    }


_JSON_FD = None


def _claim_stdout():
    """The contract is ONE JSON line on stdout.  Libraries (NCCL's version banner, torchrun notes)
    write to fd 1 behind Python's back, so everything but the result line is sent to stderr."""
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)


def emit(line):
    os.write(_JSON_FD if _JSON_FD is not None else 1, (json.dumps(line) + "\n").encode())


def main():
    args = parse_args()
    _claim_stdout()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    cores = PHYS_CORES

    if args.impl == "reference":
        if rank != 0:
            return
        N = args.N
        steps, warmup = args.steps, args.warmup
        model = build_model(N)
        res = run_cpu_port(model, steps, warmup, budget_s=float(os.environ.get("HYB_REF_BUDGET_S", 150)))
        line = {
            "impl": "reference", "metric": "ipm_iterations_per_second", "value": res["iters_per_s"], "unit": "iterations/s",
            "n_gpus": args.gpus, "steps": res["steps"], "warmup": warmup, "ms_per_step": res["ms_per_step"],
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(N, model),
            "cpu_baseline": {"value": res["iters_per_s"], "unit": "iterations/s", "cores": res["threads"], "kind": "port",
                             "physical_cores": PHYS_CORES, "logical_cpus": LOGICAL_CPUS, "blas_threads": res["threads"],
                             "sample": f"{res['steps']} interior-point iterations of the same model after {warmup} warm-up "
                                       f"(CPU restatement of the reference, numpy/scipy OpenBLAS with {res['threads']} threads, bounded "
                                       f"to ~150 s of iterations; Julia is not available on this box)",
                             "schur_assemble_ms": res["assemble_ms"], "cholesky_ms": res["factor_ms"]},
            "e2e": {"value": res["iters_per_s"], "unit": "iterations/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        }
        emit(line)
        return

    os.environ.setdefault("NCCL_DEBUG", "WARN")  # keep NCCL's version banner off stdout (one JSON line only)
    if os.environ.get("NCCL_DEBUG", "").upper() == "VERSION":
        os.environ["NCCL_DEBUG"] = "WARN"
    import torch

    # Host BLAS threads: the (untimed) preprocessing is LAPACK-heavy and runs on rank 0 only (it
    # broadcasts its outcome), so rank 0 gets all cores for the set-up; the iteration loop does
    # only O(n+q) vector work on the host, where a large BLAS pool just adds wake-up latency
    # (and oversubscribes the cores when several ranks share the box).
    try:
        from threadpoolctl import threadpool_limits

        threadpool_limits(cores if rank == 0 else 1)
        loop_threads = max(1, min(4, cores // max(world, 1)))
    except Exception:
        threadpool_limits, loop_threads = None, None
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the hypatia_b200 hot path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist_mod

        dist = dist_mod
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    import hypatia_b200  # noqa: F401
    from hypatia_b200.gpu import make_gpu_solver

    N = args.N
    model = build_model(N)
    steps, warmup = args.steps, max(args.warmup, 3)
    # multi-GPU form: "rootworker" (default; one process per GPU, rootworker.py), "replicated" (sharded.py) or
    # "single_process" (rank 0 drives every GPU through hyb_multi_* alone -- the form a Julia host would use;
    # the other ranks only keep the launch contract: they join the final barrier and reductions)
    mode = os.environ.get("HYB_MULTI_MODE", "single_process")
    single = world > 1 and mode == "single_process"
    if single and rank != 0:
        torch.cuda.synchronize()
        dist.barrier()
        for _ in range(2):
            z = torch.zeros(1, dtype=torch.float64, device="cuda")
            dist.all_reduce(z, op=dist.ReduceOp.MAX)
        dist.destroy_process_group()
        return
    if single:
        gpu = make_gpu_solver(0, iter_limit=warmup + steps, **BENCH_TOL, devices=list(range(world)))
    else:
        gpu = make_gpu_solver(local_rank, iter_limit=warmup + steps, **BENCH_TOL, dist=dist,
                              use_peer=not os.environ.get("HYB_NO_PEER"), mode=mode)
    rootworker = world > 1 and mode == "rootworker"
    if rootworker and rank != 0:
        # assembly worker: it has no iteration loop of its own; it serves the root until the root stops it,
        # once per solve of the root (the root says after each solve whether another one follows)
        while True:
            gpu.load(model)
            gpu.solve()
            again = torch.zeros(1, dtype=torch.float64, device="cuda")
            dist.broadcast(again, src=0)
            if again.item() < 0.5:
                break
        torch.cuda.synchronize()
        dist.barrier()
        for _ in range(2):  # the root's max-over-ranks reductions of wall and device time
            z = torch.zeros(1, dtype=torch.float64, device="cuda")
            dist.all_reduce(z, op=dist.ReduceOp.MAX)
        dist.destroy_process_group()
        return
    ctx = gpu.syssolver.ctx

    def sync():
        torch.cuda.synchronize()

    timer = IterTimer(ctx, sync)
    sampler = ClockSampler(local_rank)
    state = {"sampling": False}

    def callback(solver):
        if dist is not None and not rootworker and not single:
            dist.barrier()  # replicated form: every rank brackets its own iteration (root/worker: the
            # workers are inside every iteration's barrier already, only the root has an iteration loop)
        timer(solver)
        if len(timer.rows) >= warmup and not state["sampling"]:  # timed region starts after the warm-up iterations
            state["sampling"] = True
            sampler.start()
            timer._t = time.perf_counter()
        if len(timer.rows) >= warmup + steps:
            solver.iter_limit = solver.num_iters  # exactly `steps` timed iterations

    gpu.iter_callback = callback
    orig_init = gpu._initialize_point

    def init_and_start():
        orig_init()
        if threadpool_limits is not None:
            threadpool_limits(loop_threads)
        if solves and not rootworker:
            # A later solve starts after a second or so of host-only set-up with idle GPUs (and idle NVLink): its
            # first iterations would run 2-6 ms slower than the steady state the W warm-up iterations of the first
            # solve are there to reach.  Re-warm, untimed, with what iteration 0 does first: update_lhs at the
            # initial point (assembly on every GPU + factorization), twice; iteration 0 recomputes it.
            for _ in range(2):
                gpu.syssolver.update_lhs(gpu)
        timer.start()

    gpu._initialize_point = init_and_start
    if rootworker:
        timer.shard_ctx = lambda: gpu.syssolver.shard.ctx
    # The model converges in fewer iterations than warm-up + steps may ask for: the timed iterations
    # then continue into another solve of the same model (its set-up is outside the timed iterations,
    # like the first solve's), until exactly `steps` iterations have been timed.
    solves = []
    setup_s = {}
    while True:
        if threadpool_limits is not None:
            threadpool_limits(cores if rank == 0 else 1)
        gpu.iter_limit = max(1, warmup + steps - len(timer.rows))
        if rootworker or not solves:
            gpu.load(model)
        else:
            # another solve of the same model through the reference's warm re-solve path (Solvers.modify_c!,
            # Solvers.jl:296-311, then setup_modified :431-452): the system solver stays loaded, the initial point
            # is rebuilt -- no device buffer is freed and reallocated between two timed iterations
            gpu.modify_c(model.c)
        gpu.solve()
        sync()
        solves.append({"status": gpu.status, "iterations": int(gpu.num_iters)})
        if not setup_s:
            setup_s = {name: float(getattr(gpu, "time_" + name, 0.0)) for name in ("rescale", "inity", "initx", "loadsys")}
        more = len(timer.rows) < warmup + steps and len(solves) < 8
        if rootworker:
            dist.broadcast(torch.tensor([1.0 if more else 0.0], dtype=torch.float64, device="cuda"), src=0)
        if not more:
            break
    if dist is not None:
        dist.barrier()
    clocks = sampler.stop()
    rows = timer.rows[warmup:warmup + steps]
    if not rows:
        raise SystemExit(f"bench.py: the solves ended after {len(timer.rows)} iterations, before the timed region")
    k = len(rows)
    wall = sum(r["wall_s"] for r in rows)  # every iteration is bracketed by a device synchronize
    if dist is not None:
        tw = torch.tensor([wall], dtype=torch.float64, device="cuda")
        dist.all_reduce(tw, op=dist.ReduceOp.MAX)
        wall = float(tw.item())
    dev_s = sum(r["dev_ms"] + r.get("reduce_ms", 0.0) for r in rows) * 1e-3
    reduce_ms = float(np.mean([r.get("reduce_ms", 0.0) for r in rows]))
    if dist is not None:
        td = torch.tensor([dev_s], dtype=torch.float64, device="cuda")
        dist.all_reduce(td, op=dist.ReduceOp.MAX)
        dev_s = float(td.item())
    ms_per_step = 1e3 * wall / k
    h2d = sum(r["h2d"] for r in rows) / k
    d2h = sum(r["d2h"] for r in rows) / k
    t_syrk = float(np.mean([r["t_syrk"] for r in rows]))
    t_w = float(np.mean([r["t_w"] for r in rows]))
    t_fact = float(np.mean([r["t_fact"] for r in rows]))
    launches = int(sum(r["launches"] for r in rows))
    if rank != 0:
        return

    # every cone of this model takes the sqrt-oracle (syrk) branch, qrchol.jl:228-242; with N ranks the
    # launch timed on rank 0 contracts that rank's rows only
    if rootworker:
        q_sqrt = gpu.syssolver.row_counts[0]
    elif single:
        q_sqrt = max(hi - lo for lo, hi in gpu.syssolver.row_ranges)
    else:
        q_sqrt = gpu.cones.row_counts[0] if world > 1 else model.q
    syrk_flops = q_sqrt * N * (N + 1)  # algorithmic flops of the G'HG contraction per launch
    peak_tf = cublas_dgemm_tflops()
    achieved_tf = syrk_flops / (t_syrk * 1e-3) / 1e12
    line = {
        "metric": "ipm_iterations_per_second",
        "value": k / dev_s if dev_s > 0 else None,
        "unit": "iterations/s",
        "n_gpus": world,
        "steps": k,
        "warmup": warmup,
        "ms_per_step": ms_per_step,
        "higher_is_better": True,
        "scaling": "strong",
        "vs_baseline": None,
        "dtype": "f64",
        "data": "synthetic",
        "config": workload_config(N, model, {
            "value_definition": "iterations / device time: CUDA-event time of every kernel region of the C-ABI calls of the timed iterations (inputs resident in HBM)",
            "e2e_definition": "iterations / wall time of the same iterations through the C-ABI with host vectors (all H2D/D2H copies and the host-side stepper included)",
            "solves": solves,
            "parallelism": ("cones partitioned by block over %d ranks; partial Schur matrices reach rank 0 by %s; Cholesky + solves on rank 0"
                            % (world, "peer stores from the syrk epilogue over NVLink + fixed-order sum" if (rootworker or single or getattr(gpu.syssolver.backend, "peer", False)) else "NCCL reduce")) if world > 1 else "single GPU",
            "multi_gpu_form": ("root/worker: rank 0 runs the algorithm, 1 point broadcast + 1 barrier per iteration" if rootworker else
                               ("single host process driving all GPUs through hyb_multi_* (peer access, no collective library)" if single
                                else "replicated host driver, sharded oracles")) if world > 1 else None,
            "line_search_candidates_per_step": float(np.mean([r["n_cand"] for r in rows])),
            "refinement_solves_per_step": float(np.mean([r["n_ref"] for r in rows])),
        }),
        "e2e": {"value": k / wall, "unit": "iterations/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
        "gpu_launches": launches,
        # untimed set-up of the solve (Solvers.jl:395-429), seconds: rescale, primal-equality QR + reduce,
        # dual-equality pivoted QR (hyb_geqp3 on the GPU) + initial point, upload of G*Q2
        "setup_s": setup_s,
        "phase_ms": {name: float(np.mean([r["ph_" + name] for r in rows])) for name in ("upsys", "uprhs", "getdir", "search")},
        # spread of the timed iterations (wall, ms): a host-side stall in one iteration shows here, not in the mean
        "step_ms_min_median_max": [float(1e3 * f([r["wall_s"] for r in rows])) for f in (np.min, np.median, np.max)],
        "step_ms": [round(1e3 * r["wall_s"], 2) for r in rows],
        "step_upsys_ms": [round(r["ph_upsys"], 2) for r in rows],
        "schur_assemble_ms": t_w + t_syrk,
        "schur_reduce_ms": reduce_ms,
        "cholesky_ms": t_fact,
        "clocks": clocks,
        "roofline": {
            "kernel": "gemm_tn_kernel (G'HG syrk, FP64 DMMA, TMA-fed)",
            "bound": "tensor", "achieved": achieved_tf, "peak": peak_tf, "unit": "TFLOP/s",
            "frac": achieved_tf / peak_tf if peak_tf else None,
            "peak_source": "cuBLAS DGEMM 8192^3 measured in this run (MEASURED_PEAKS.json has no FP64 figure; FP64 tensor work on sm_100a is DMMA, not tcgen05)",
            "algorithmic_flops_per_launch": syrk_flops, "avg_launch_ms": t_syrk,
            "traffic": ncu_traffic(N, q_sqrt),
            "traffic_source": "dram__bytes_read.sum + dram__bytes_write.sum of one launch, ncu --set full capture in profiles/r02_syrk_N8192_ncu_full.json (same N, K)",
        },
    }
    if not args.no_cpu_baseline and world == 1:
        steps_cpu = 3
        res = run_cpu_port(build_model(N), steps_cpu, 1, preset_fact=getattr(gpu, "AG_fact", None), budget_s=30.0)
        line["cpu_baseline"] = {
            "value": res["iters_per_s"], "unit": "iterations/s", "cores": res["threads"], "kind": "port",
            "physical_cores": PHYS_CORES, "logical_cpus": LOGICAL_CPUS, "blas_threads": res["threads"],
            "sample": f"{res['steps']} interior-point iterations of the same model after 1 warm-up iteration "
                      f"(CPU restatement of the reference, numpy/scipy OpenBLAS, {res['threads']} BLAS threads on "
                      f"{PHYS_CORES} physical cores)",
            "schur_assemble_ms": res["assemble_ms"], "cholesky_ms": res["factor_ms"],
        }
    emit(line)
    if dist is not None:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
