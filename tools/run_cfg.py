"""Run one BASELINE config on the GPU path (and optionally the CPU oracle) and print timing / parity."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import hypatia_b200
from hypatia_b200 import instances
from hypatia_b200.gpu import make_gpu_solver
from hypatia_b200.host.solver import Solver

name = sys.argv[1]
with_cpu = len(sys.argv) > 2 and sys.argv[2] == "cpu"
builders = {
    "cfg1": lambda: instances.linearopt(500, 1000, seed=1),
    "cfg2": lambda: instances.nearestpsd(side=300, seed=1),
    "cfg2s": lambda: instances.nearestpsd(side=100, seed=1),
    "cfg3": lambda: instances.polymin(n=2, halfdeg=30, seed=1),
    "cfg3s": lambda: instances.polymin(n=2, halfdeg=12, seed=1),
    "cfg4": lambda: instances.portfolio_socs(K=100_000, d=4, n=2048, seed=1),
    "cfg4s": lambda: instances.portfolio_socs(K=20_000, d=4, n=512, seed=1),
    "cfg4n": lambda: instances.portfolio_norminf(K=100_000, d=4, n=2048, seed=1),
    "cfg4ns": lambda: instances.portfolio_norminf(K=20_000, d=4, n=512, seed=1),
}
# cfg4: the reference's benchmark tolerances (benchmarks/natvsext/run.jl:44-60)
kw = dict(tol_rel_opt=1e-7, tol_feas=1e-7, tol_abs_opt=1e-10, tol_infeas=1e-10) if name.startswith("cfg4") else {}
t0 = time.time(); model = builders[name](); tb = time.time() - t0
gpu = make_gpu_solver(0, trace=True, **kw)
gpu.load(model)
t0 = time.time(); gpu.solve(); tg = time.time() - t0
ss = gpu.syssolver
out = dict(config=name, q=int(model.q), n=int(gpu.model.n), build_s=tb, gpu_status=gpu.status, gpu_iters=gpu.num_iters, gpu_obj=gpu.primal_obj,
           gpu_solve_s=tg, gpu_setup_s=gpu.time_rescale + gpu.time_initx + gpu.time_inity + gpu.time_loadsys,
           gpu_iter_loop_s=gpu.time_upsys + gpu.time_uprhs + gpu.time_getdir + gpu.time_search,
           t_upsys=gpu.time_upsys, t_getdir=gpu.time_getdir, t_search=gpu.time_search, t_uprhs=gpu.time_uprhs,
           assemble_ms_per_iter=ss.time_assemble_ms / max(ss.num_updates, 1), fact_ms_per_iter=ss.time_fact_ms / max(ss.num_updates, 1),
           launches=ss.ctx.launch_count())
if with_cpu:
    from oracle.cones import CpuConeSet
    from oracle.qrchol import QRCholDenseSystemSolver
    cpu = Solver(CpuConeSet, QRCholDenseSystemSolver(), trace=True, **kw)
    cpu.load(builders[name]())
    t0 = time.time(); cpu.solve(); tc = time.time() - t0
    devs = []
    for a, b in zip(gpu.trace, cpu.trace):
        devs.append(max(abs(a[k] - b[k]) / max(abs(b[k]), 1e-6 if k.endswith("_res") else 1e-12) for k in ("x_norm_res", "z_norm_res", "mu", "gap", "tau")))
    first_bad = next((i for i, d in enumerate(devs) if d > 1e-8), None)
    out.update(cpu_status=cpu.status, cpu_iters=cpu.num_iters, cpu_obj=cpu.primal_obj, cpu_solve_s=tc,
               cpu_iter_loop_s=cpu.time_upsys + cpu.time_uprhs + cpu.time_getdir + cpu.time_search,
               obj_rel_diff=abs(gpu.primal_obj - cpu.primal_obj) / (1 + abs(cpu.primal_obj)),
               first_iter_traj_dev_gt_1e8=first_bad, traj_dev=[float(f"{d:.2e}") for d in devs])
print(json.dumps(out))
os.makedirs("gpurun_out", exist_ok=True)
json.dump(out, open(f"gpurun_out/run_{name}.json", "w"), indent=1)
