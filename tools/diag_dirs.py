"""Compare batched (device) and unbatched (host pattern) direction solves on one instance."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import hypatia_b200  # noqa
from hypatia_b200 import instances
from hypatia_b200.gpu import make_gpu_solver
from hypatia_b200.host.solver import Solver
from oracle.cones import CpuConeSet
from oracle.qrchol import QRCholDenseSystemSolver

mk = lambda: instances.linearopt(m=60, n=200, seed=3)
runs = {}
for name, kw in (("batched", dict(batched_dirs=True)), ("unbatched", dict(batched_dirs=False))):
    s = make_gpu_solver(0, trace=True, **kw)
    log = []
    s.iter_callback = lambda sv, log=log: log.append((sv.stepper.last_num_ref, sv.stepper.last_num_cand, sv.worst_dir_res))
    s.load(mk()); s.solve()
    runs[name] = (s, log)
    print(name, s.status, s.num_iters, s.primal_obj)
ref = Solver(CpuConeSet, QRCholDenseSystemSolver(), trace=True)
log = []
ref.iter_callback = lambda sv, log=log: log.append((sv.stepper.last_num_ref, sv.stepper.last_num_cand, sv.worst_dir_res))
ref.load(mk()); ref.solve()
print("oracle", ref.status, ref.num_iters, ref.primal_obj)
for i in range(max(len(runs["batched"][1]), len(log))):
    row = []
    for l in (runs["batched"][1], runs["unbatched"][1], log):
        row.append(l[i] if i < len(l) else None)
    tr = [r.trace[i]["mu"] if i < len(r.trace) else None for r in (runs["batched"][0], runs["unbatched"][0], ref)]
    print(i, row, tr)
