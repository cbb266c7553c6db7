"""Device-resident stepper vs host-vector protocol on one known-answer instance (diagnostics)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hypatia_b200
from hypatia_b200 import instances
from hypatia_b200.gpu import make_gpu_solver
name = sys.argv[1] if len(sys.argv) > 1 else "nonnegative4"
for host in (False, True):
    if host:
        os.environ["HYB_HOST_STEPPER"] = "1"
    s = make_gpu_solver(0, trace=True)
    s.load(instances.KNOWN_ANSWER[name]()[0])
    s.solve()
    print("HOST" if host else "DEVICE", s.status, s.num_iters, s.primal_obj)
    for r in s.trace:
        print("  it %2d mu %.6e xres %.3e zres %.3e tau %.6e kap %.6e alpha %.4f ncand %d nref %d dirres %.2e" % (
            r["iter"], r["mu"], r["x_norm_res"], r["z_norm_res"], r["tau"], r["kap"], r["alpha"], r["n_cand"], r["n_ref"], r["worst_dir_res"]))
