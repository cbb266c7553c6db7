import functools

import numpy as np

import hypatia_b200  # noqa: F401


@functools.lru_cache(maxsize=1)
def context():
    from hypatia_b200.lib import Context

    return Context(0)


def rel_err(a, b):
    a, b = np.asarray(a), np.asarray(b)
    denom = max(float(np.max(np.abs(b))), 1e-300) if b.size else 1.0
    return float(np.max(np.abs(a - b))) / denom if b.size else 0.0
