"""Side-by-side GPU / CPU-oracle trajectories for one instance (diagnostic)."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

import hypatia_b200  # noqa: F401
from hypatia_b200 import instances
from hypatia_b200.gpu import make_gpu_solver
from hypatia_b200.host.solver import Solver
from oracle.cones import CpuConeSet
from oracle.qrchol import QRCholDenseSystemSolver

name = sys.argv[1]
if name in instances.KNOWN_ANSWER:
    fn = lambda: instances.KNOWN_ANSWER[name]()[0]
elif name == "cfg1":
    fn = lambda: instances.linearopt(500, 1000, seed=1)
elif name.startswith("block"):
    fn = lambda: instances.block_multicone(N=int(name[5:]), seed=1)
gpu = make_gpu_solver(0, trace=True, batched_dirs=not os.environ.get("HYB_UNBATCHED"))
cpu = Solver(CpuConeSet, QRCholDenseSystemSolver(), trace=True)
for s in (gpu, cpu):
    s.load(fn())
    s.solve()
print("status", gpu.status, cpu.status, "iters", gpu.num_iters, cpu.num_iters, "obj", gpu.primal_obj, cpu.primal_obj)
keys = ("x_norm_res", "z_norm_res", "mu", "tau", "worst_dir_res", "alpha", "n_cand", "n_ref", "prox")
for i in range(max(len(gpu.trace), len(cpu.trace))):
    a = gpu.trace[i] if i < len(gpu.trace) else None
    b = cpu.trace[i] if i < len(cpu.trace) else None
    for tag, r in (("G", a), ("C", b)):
        if r is not None:
            print(i, tag, " ".join(f"{k}={r[k]:.6e}" if isinstance(r[k], float) else f"{k}={r[k]}" for k in keys))
print("fact kinds / updates:", gpu.syssolver.last_fact_kind, gpu.syssolver.num_updates)
