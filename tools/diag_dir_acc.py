"""Where does the direction accuracy go?  GPU vs oracle components at the oracle's iterate (diagnostics)."""
import os, sys
os.environ["HYB_HOST_STEPPER"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import test_gpu_dirs as t
from gpu_util import rel_err
from hypatia_b200.host.solver import Solver
from hypatia_b200.gpu import make_gpu_solver
from oracle.cones import CpuConeSet
from oracle.qrchol import QRCholDenseSystemSolver
name, n_iters = sys.argv[1], int(sys.argv[2])
ref = Solver(CpuConeSet, QRCholDenseSystemSolver(), iter_limit=n_iters); ref.load(t.MODELS[name]()); ref.solve()
pv = ref.point.vec.copy()
cpu = Solver(CpuConeSet, QRCholDenseSystemSolver()); t._at_point(cpu, t.MODELS[name], pv)
gpu = make_gpu_solver(0); t._at_point(gpu, t.MODELS[name], pv)
rng = np.random.default_rng(0)
q, n = cpu.model.q, cpu.model.n
v = rng.standard_normal(q)
offs = cpu.cones.offs
def per_block(a, b):
    out = {}
    for k, s in enumerate(cpu.cones.specs):
        sl = slice(int(offs[k]), int(offs[k+1]))
        key = type(s).__name__
        e = np.max(np.abs(a[sl]-b[sl]))/max(np.max(np.abs(b[sl])),1e-300)
        out[key] = max(out.get(key,0), e)
    return out
print("mu", cpu.mu)
for fn in ("grad",):
    print(fn, per_block(gpu.cones.grad(), cpu.cones.grad()))
for fn in ("hess_prod", "inv_hess_prod", "dder3"):
    print(fn, per_block(getattr(gpu.cones, fn)(v), getattr(cpu.cones, fn)(v)))
Sg = np.triu(gpu.syssolver.schur_matrix()); Sc = np.triu(cpu.syssolver.lhs)
print("schur rel err", rel_err(Sg, Sc), "cond", np.linalg.cond(Sc + np.triu(Sc,1).T))
b = rng.standard_normal(n)
xg = t_x = None
import ctypes as C
from hypatia_b200.lib import dptr
xg = np.empty(n); gpu.syssolver.ctx.check(gpu.syssolver.lib.hyb_schur_solve(gpu.syssolver.ctx.ptr, dptr(b), dptr(xg), 1), "solve")
xc = cpu.syssolver.fact.solve(b)
S = Sc + np.triu(Sc,1).T
print("solve resid gpu", np.max(np.abs(S@xg-b)), "cpu", np.max(np.abs(S@xc-b)), "diff", rel_err(xg, xc))
print("G mul", rel_err(gpu.G_mul(xc), cpu.G_mul(xc)), rel_err(gpu.Gt_mul(v), cpu.Gt_mul(v)))
