"""Generate tests/golden/*.json (run from the repo root: python tests/golden/make_golden.py).

known_answer.json   the expected results of the reference's RNG-free known-answer instances
                    (test/nativeinstances.jl, line ranges in SURVEY.md §8c), as transcribed in
                    hypatia_b200/instances.py: status, objective, solution entries.  The reference is
                    Julia and cannot run in this image, so these numbers -- written by the reference's
                    authors into their own tests -- are the golden vectors that pin the oracle.
trajectories.json   what the pinned CPU oracle produces on those instances and on a few seeded
                    synthetic ones: status, iteration count, objective, mu and step length per iteration.
                    The CUDA path is compared with these committed numbers on the GPU box
                    (tests/test_golden.py), where /root/reference does not exist.
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import hypatia_b200  # noqa: E402,F401
from hypatia_b200 import instances  # noqa: E402
from hypatia_b200.host.solver import Solver  # noqa: E402
from oracle.cones import CpuConeSet  # noqa: E402
from oracle.qrchol import QRCholDenseSystemSolver  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))

SEEDED = {
    "block_multicone_N64": ("block_multicone", dict(N=64, seed=0, psd_side=8)),
    "block_multicone_N160": ("block_multicone", dict(N=160, seed=5, psd_side=8)),
    "linearopt_60x200": ("linearopt", dict(m=60, n=200, seed=3)),
    "nearestpsd_side20": ("nearestpsd", dict(side=20, seed=1)),
    "polymin_n2_hd4": ("polymin", dict(n=2, halfdeg=4, seed=1)),
    "nearestdnn_side12": ("nearestdnn", dict(side=12, seed=1)),
    "sparsepsd_maxeig_side20": ("sparsepsd_maxeig", dict(side=20, seed=1)),
    "polynormeucl_n2_hd2_R3": ("polynormeucl", dict(n=2, halfdeg=2, R=3, seed=1)),
    "polynormone_n2_hd2_R3": ("polynormone", dict(n=2, halfdeg=2, R=3, seed=1)),
    # the same families at sizes where the stored Hessian is a few hundred rows
    "nearestdnn_side30": ("nearestdnn", dict(side=30, seed=2)),
    "sparsepsd_maxeig_side80": ("sparsepsd_maxeig", dict(side=80, seed=2)),
    "polynormeucl_n2_hd5_R3": ("polynormeucl", dict(n=2, halfdeg=5, R=3, seed=2)),
    "polynormone_n2_hd5_R4": ("polynormone", dict(n=2, halfdeg=5, R=4, seed=2)),
}


def _jsonable(v):
    if isinstance(v, dict):
        return {str(k): _jsonable(x) for k, x in v.items()}
    if isinstance(v, (np.ndarray, list, tuple)):
        return [float(x) for x in np.asarray(v, dtype=np.float64).ravel()]
    if isinstance(v, (np.floating, float, int, np.integer)):
        return float(v)
    return v


def build(name):
    if name in instances.KNOWN_ANSWER:
        return instances.KNOWN_ANSWER[name]()[0]
    fn, kw = SEEDED[name]
    return getattr(instances, fn)(**kw)


def main():
    known = {name: _jsonable(instances.KNOWN_ANSWER[name]()[1]) for name in sorted(instances.KNOWN_ANSWER)}
    json.dump(known, open(os.path.join(HERE, "known_answer.json"), "w"), indent=1, sort_keys=True)
    traj = {}
    for name in sorted(instances.KNOWN_ANSWER) + sorted(SEEDED):
        s = Solver(CpuConeSet, QRCholDenseSystemSolver(), trace=True)
        s.load(build(name))
        s.solve()
        traj[name] = dict(
            status=s.status,
            num_iters=int(s.num_iters),
            primal_obj=float(s.primal_obj),
            dual_obj=float(s.dual_obj),
            mu=[float(r["mu"]) for r in s.trace],
            alpha=[float(r["alpha"]) for r in s.trace],
        )
        print(name, s.status, s.num_iters, s.primal_obj)
    json.dump(traj, open(os.path.join(HERE, "trajectories.json"), "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
