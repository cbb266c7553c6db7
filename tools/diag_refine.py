"""Print the refinement rounds of the last iteration for the device (batched) and host patterns."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import hypatia_b200  # noqa
from hypatia_b200 import instances
from hypatia_b200.gpu import make_gpu_solver
name, iters = sys.argv[1], int(sys.argv[2])
batched = not os.environ.get("HYB_UNBATCHED")
s = make_gpu_solver(0, trace=True, iter_limit=iters, batched_dirs=batched)
def cb(sv):
    print(f"--- end of iteration {sv.num_iters} (n_ref {sv.stepper.last_num_ref}, worst {sv.worst_dir_res:.3e})", file=sys.stderr, flush=True)
s.iter_callback = cb
s.load(instances.KNOWN_ANSWER[name]()[0]); s.solve()
