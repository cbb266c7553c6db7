"""At a late iteration of an instance: device-reported residual vs host-evaluated residual of the
same direction, with and without refinement."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import hypatia_b200  # noqa
from hypatia_b200 import instances
from hypatia_b200.gpu import make_gpu_solver
from hypatia_b200.host.models import Point
from hypatia_b200.host.syscommon import apply_lhs, get_directions
from hypatia_b200.host import stepper as st
name, iters = sys.argv[1], int(sys.argv[2])
s = make_gpu_solver(0, iter_limit=iters)
s.load(instances.KNOWN_ANSWER[name]()[0]); s.solve()
ss, stp, m = s.syssolver, s.stepper, s.model
ss.update_lhs(s)
r1, r2, d1, d2 = (Point(m.n, m.p, m.q) for _ in range(4))
st.update_rhs_cent(s, r1); st.update_rhs_pred(s, r2)
def host_res(dv, rv):
    stp.dir.vec[:] = dv
    res = apply_lhs(stp, s)
    return res.vec - rv
for max_ref in (0, 5):
    s.max_ref_steps = max_ref
    ss._dev_const_stale = True
    ss.get_directions_batch(stp, s, (r1, r2), (d1, d2))
    print("max_ref", max_ref, "device-reported", ss.last_res_norms, ss.last_num_refs)
    for d, r in ((d1, r1), (d2, r2)):
        hr = host_res(d.vec.copy(), r.vec)
        n, q = m.n, m.q
        print("   host-evaluated inf-norm %.3e | x %.2e z %.2e tau %.2e s %.2e kap %.2e" % (
            np.abs(hr).max(), np.abs(hr[:n]).max(), np.abs(hr[n:n+q]).max(), abs(hr[n+q]), np.abs(hr[n+q+1:n+2*q+1]).max(), abs(hr[-1])))
s.max_ref_steps = 5
ss._host_const_stale = True
for r in (r1, r2):
    stp.rhs.vec[:] = r.vec
    stp.last_num_ref = 0
    get_directions(stp, s)
    hr = host_res(stp.dir.vec.copy(), r.vec)
    print("host pattern: refs", stp.last_num_ref, "res %.3e" % np.abs(hr).max())
