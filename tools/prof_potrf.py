import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hypatia_b200
from hypatia_b200.lib import Context
N = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
ctx = Context(0)
mf, ms = ctx.bench_potrf(N, reps=1)
print(f"potrf N={N}: {mf:.2f} ms; potrs {ms:.3f} ms")
