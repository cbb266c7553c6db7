"""ORACLE — TEST INFRASTRUCTURE ONLY.  CPU restatement of the reference hot
path (cone oracles, Schur assembly, dense Cholesky + solves).  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import it; the product path (hypatia.jl_b200/gpu.py + csrc/) never does."""
import os as _os
import sys as _sys

_root = _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__)))
if _root not in _sys.path:
    _sys.path.insert(0, _root)
