"""cProfile of the host-side set-up (rescale, primal / dual equality handling, initial point, model upload) of one
BASELINE configuration on the GPU path: python tools/prof_setup.py cfg4"""
import cProfile
import os
import pstats
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hypatia_b200  # noqa: F401
from hypatia_b200 import instances
from hypatia_b200.gpu import make_gpu_solver

name = sys.argv[1] if len(sys.argv) > 1 else "cfg4s"
build = {"cfg4s": lambda: instances.portfolio_socs(K=20_000, d=4, n=512, seed=1),
         "cfg4": lambda: instances.portfolio_socs(K=100_000, d=4, n=2048, seed=1),
         "cfg2": lambda: instances.nearestpsd(side=300, seed=1)}[name]
t0 = time.time()
m = build()
print("build", round(time.time() - t0, 2), "q", m.q, "n", m.n, "p", m.p)
s = make_gpu_solver(0)
s.load(m)
pr = cProfile.Profile()
pr.enable()
s._setup_solver()
s._setup_loaded()
pr.disable()
print({k: round(getattr(s, "time_" + k), 3) for k in ("rescale", "initx", "inity", "loadsys")})
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
