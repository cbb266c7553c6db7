"""Actual GPU-vs-oracle errors of the WSOS cone oracles on the test sets (to set honest tolerances)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import test_gpu_cones as t
from gpu_util import rel_err
for name in ("wsos", "wsospsd"):
    specs = t.SPEC_SETS[name]
    rng = np.random.default_rng(sorted(t.SPEC_SETS).index(name) + 100)
    cpu, gpu = t.make_pair(specs)
    primal = t.interior_point(specs, rng, scale=0.7); dual = t.interior_point(specs, rng, scale=1.3)
    for c in (cpu, gpu):
        c.load_point(primal, 1.1); c.load_dual_point(dual); c.reset_data()
    assert cpu.is_feas() and gpu.is_feas() and cpu.is_dual_feas()
    print(name, "grad", rel_err(gpu.grad(), cpu.grad()))
    for um in (True, False):
        a, b = cpu.check_neighborhood(1.1, um, 0.01, 1e9), gpu.check_neighborhood(1.1, um, 0.01, 1e9)
        print(name, "agg", um, a, b, abs(a[1]-b[1])/max(1,abs(a[1])))
    for m in (1, 7):
        arr = rng.standard_normal(cpu.q) if m == 1 else np.asfortranarray(rng.standard_normal((cpu.q, m)))
        for fn in ("hess_prod", "inv_hess_prod", "block_hess_prod", "hess_prod_slow", "sqrt_hess_prod", "inv_sqrt_hess_prod"):
            try:
                want = getattr(cpu, fn)(arr) if hasattr(cpu, fn) else None
            except Exception:
                want = None
            if want is None:
                want = np.empty_like(arr)
                for k, c in enumerate(cpu.cones):
                    sl = slice(int(cpu.offs[k]), int(cpu.offs[k + 1]))
                    c.use_sqrt_hess_oracles(c.dim + 1)
                    want[sl] = getattr(c, fn)(arr[sl])
            print(name, fn, m, rel_err(getattr(gpu, fn)(arr), want))
    d = 0.1 * rng.standard_normal(cpu.q)
    print(name, "dder3", rel_err(gpu.dder3(d), cpu.dder3(d)))
    # conditioning of the Hessians involved
    for c in cpu.cones:
        if hasattr(c, "hess_mat") or True:
            try:
                print("   cone dim", c.dim, "cond(H) %.2e" % np.linalg.cond(c.hess()))
            except Exception as e:
                pass
