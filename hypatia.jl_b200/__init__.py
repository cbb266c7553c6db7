"""hypatia_b200 — B200-native interior-point hot path behind Hypatia.jl's own
SystemSolver / Cone interfaces.

Layout
  csrc/    CUDA kernels (sm_100a) + the C-ABI (include/hypatia_b200.h)
  lib.py   ctypes binding of the C-ABI (fails loudly if the .so is missing)
  host/    host-side mirror of the reference's Solver / Stepper / search /
           SystemSolver plumbing (stands in for the Julia host, which is absent
           from this image); generic over a cone set and a system solver
  gpu.py   B200QRCholSystemSolver + GPU-resident cone batch (the product path)
  cones_spec.py, polyutils.py, instances.py   model description / builders
"""
from . import cones_spec  # noqa: F401
