"""Model builders: the reference's RNG-free known-answer instances (used as golden
vectors) and the synthetic workloads of BASELINE.json configs 1-5.

Known-answer instances restate test/nativeinstances.jl (line ranges cited per
function); they return (model, expected) where ``expected`` holds the status
and the values the reference test asserts.  Config builders follow
SURVEY.md §8d (numpy default_rng(seed); Julia's RNG stream is not reproducible).
"""
from __future__ import annotations

import math
from typing import Dict, Tuple

import numpy as np

from . import cones_spec as cs
from . import polyutils as pu
from .host.models import Model

RT2 = math.sqrt(2.0)
RT3 = math.sqrt(3.0)


def _I(n):
    return np.eye(n)


# ---------------------------------------------------------------------------
# known-answer instances (test/nativeinstances.jl)
# ---------------------------------------------------------------------------
def dimension1():  # :108-128
    m = Model([-1, 0], np.zeros((0, 2)), [], [[1, 0]], [1], [cs.Nonnegative(1)])
    return m, dict(status="Optimal", primal_obj=-1.0, x=[1, 0])


def primalinfeas1():  # :189-200
    m = Model([1, 0], [[1, 1]], [-2], -_I(2), [0, 0], [cs.Nonnegative(2)])
    return m, dict(status="PrimalInfeasible")


def primalinfeas2():  # :202-213
    G = np.vstack([-_I(3), np.diag([1.0, 1.0, -1.0])])
    h = [0, 0, 0, 1, 1, -2]
    m = Model([1, 1, 1], np.zeros((0, 3)), [], G, h, [cs.EpiNormEucl(3), cs.Nonnegative(3)])
    return m, dict(status="PrimalInfeasible")


def dualinfeas2():  # :241-252
    m = Model([-1, 0], np.zeros((0, 2)), [], [[-1, 0], [0, 0], [0, -1]], [0, 1, 0], [cs.EpiPerSquare(3)])
    return m, dict(status="DualInfeasible")


def dualinfeas3():  # :254-265
    m = Model([0, 1, 1, 0], np.zeros((0, 4)), [], -_I(4), np.zeros(4), [cs.EpiPerSquare(4)])
    return m, dict(status="DualInfeasible")


def nonnegative4():  # :312-327
    G = np.zeros((3, 2))
    for r, c_, v in zip([1, 1, 2, 3], [1, 2, 2, 2], [1, -1, 1, -1]):
        G[r - 1, c_ - 1] = v
    m = Model([-2, 0], np.zeros((0, 2)), [], G, [0, 2, 0], [cs.Nonnegative(3)])
    return m, dict(status="Optimal", primal_obj=-4.0, x=[2, 2], s=[0, 0, 2], z=[2, 2, 0])


def possemideftri1():  # :329-342
    m = Model([0, -1, 0], [[1, 0, 0], [0, 0, 1]], [0.5, 1], -_I(3), np.zeros(3), [cs.PosSemidefTri(3)])
    return m, dict(status="Optimal", primal_obj=-1.0, x_at={1: 1.0})


def possemideftri2():  # :344-357
    m = Model([0, -1, 0], [[1, 0, 1]], [0], -_I(3), np.zeros(3), [cs.PosSemidefTri(3)])
    return m, dict(status="Optimal", primal_obj=0.0, x_norm=0.0)


def possemideftri8():  # :458-496
    G = np.zeros((15, 1))
    G[[0, 2, 5, 9, 14], 0] = -1
    h = np.zeros(15)
    h[[6, 7, 8, 10, 11, 12]] = RT2 * np.array([1, 1, 0, 1, -1, 1.0])
    m = Model([1], np.zeros((0, 1)), [], G, h, [cs.PosSemidefTri(15)])
    inv6, rt2inv6, invrt6 = 1 / 6, RT2 / 6, 1 / (RT2 * RT3)
    s = [RT3, 0, RT3, 0, 0, RT3, RT2, RT2, 0, RT3, RT2, -RT2, RT2, 0, RT3]
    z = [inv6, -rt2inv6, inv6, rt2inv6, -rt2inv6, inv6, 0, 0, 0, 0, -invrt6, invrt6, -invrt6, 0, 0.5]
    return m, dict(status="Optimal", primal_obj=RT3, s=s, z=z)


def possemideftri9():  # :498-556
    G = np.zeros((16, 10))
    for (i, j) in [(1, 2), (1, 4), (1, 7), (1, 8), (1, 10)]:
        G[i - 1, j - 1] = 0.5
    for (i, j) in [(1, 1), (2, 2), (4, 4), (7, 7), (11, 8), (16, 10)]:
        G[i - 1, j - 1] = -1
    for (i, j) in [(3, 3), (5, 5), (6, 6), (15, 9)]:
        G[i - 1, j - 1] = -RT2
    h = np.zeros(16)
    h[[7, 8, 9, 11, 12, 13]] = RT2 * np.array([1, 1, 0, 1, -1, 1.0])
    c = np.concatenate([[1.0], np.zeros(9)])
    m = Model(c, np.zeros((0, 10)), [], G, h, [cs.Nonnegative(1), cs.PosSemidefTri(15)])
    invrt2, invrt3 = 1 / RT2, 1 / RT3
    invrt6 = invrt2 * invrt3
    s = [0, invrt2 + invrt3, 1 - RT2 / RT3, invrt2 + invrt3, RT2 * invrt3, -RT2 * invrt3, invrt3,
         RT2, RT2, 0, RT2, RT2, -RT2, RT2, 0, RT3]
    z = [1, 0.5, 0, 0.5, 0, 0, 0.5, -0.5, -0.5, 0, 0.5, -invrt6, invrt6, -invrt6, 0, 0.5]
    return m, dict(status="Optimal", primal_obj=RT2 + RT3, s=s, z=z)


def epinorminf1():  # :882-897
    m = Model([0, -1, -1], [[1, 0, 0], [0, 1, 0]], [1, 1 / RT2], -_I(3), np.zeros(3), [cs.EpiNormInf(3)])
    return m, dict(status="Optimal", primal_obj=-1 - 1 / RT2, x=[1, 1 / RT2, 1], y=[1, 1])


def epinorminf2():  # :899-928  (l = 3; a dual-barrier and a primal-barrier cone)
    l = 3
    L = 2 * l + 1
    c = np.arange(-l, l + 1, dtype=np.float64)
    A = np.zeros((2, L))
    A[0, 0] = A[0, L - 1] = A[1, 0] = 1
    A[1, L - 1] = -1
    G = np.vstack([np.zeros((1, L)), _I(L), np.zeros((1, L)), 2 * _I(L)])
    h = np.zeros(2 * L + 2)
    h[0] = 1
    h[L + 1] = 1
    cones = [cs.EpiNormInf(L + 1, use_dual=True), cs.EpiNormInf(L + 1, use_dual=False)]
    m = Model(c, A, [0, 0], G, h, cones, obj_offset=1.0)
    return m, dict(status="Optimal", primal_obj=-l + 2.0, x_at={1: 0.5, L - 2: -0.5}, x_abs_sum=1.0, tol_factor=10.0)


def epinorminf3():  # :930-946 (primal-barrier case of the loop)
    m = Model([1, 0, 0, 0, 0, 0], np.zeros((0, 6)), [], -_I(6), np.zeros(6), [cs.EpiNormInf(6)])
    return m, dict(status="Optimal", primal_obj=0.0, x_norm=0.0)


def epinorminf3_dual():  # :930-946 (use_dual = true case of the loop)
    m = Model([1, 0, 0, 0, 0, 0], np.zeros((0, 6)), [], -_I(6), np.zeros(6), [cs.EpiNormInf(6, use_dual=True)])
    return m, dict(status="Optimal", primal_obj=0.0, x_norm=0.0)


def epinorminf4():  # :948-962
    m = Model([0, 1, -1], [[1, 0, 0], [0, 1, 0]], [1, -0.4], -_I(3), np.zeros(3), [cs.EpiNormInf(3, use_dual=True)])
    return m, dict(status="Optimal", primal_obj=-1.0, x=[1, -0.4, 0.6], y=[1, 0])


def epinormeucl1():  # :1014-1030
    m = Model([0, -1, -1], [[10, 0, 0], [0, 10, 0]], [10, 10 / RT2], -_I(3), np.zeros(3), [cs.EpiNormEucl(3)])
    return m, dict(status="Optimal", primal_obj=-RT2, x=[1, 1 / RT2, 1 / RT2], y=[RT2 / 10, 0])


def epinormeucl2():  # :1032-1045
    m = Model([0, -1, -1], [[1, 0, 0]], [0], -_I(3), np.zeros(3), [cs.EpiNormEucl(3)])
    return m, dict(status="Optimal", primal_obj=0.0, x_norm=0.0)


def epinormeucl3():  # :1047-1060
    m = Model([1, 0, 0], [[0, 1, 0]], [1], -_I(3), np.zeros(3), [cs.EpiNormEucl(3)])
    return m, dict(status="Optimal", primal_obj=1.0, x=[1, 1, 0])


def epipersquare1():  # :1062-1076
    m = Model([0, 0, -1, -1], [[1, 0, 0, 0], [0, 1, 0, 0]], [0.5, 1], -_I(4), np.zeros(4), [cs.EpiPerSquare(4)])
    return m, dict(status="Optimal", primal_obj=-RT2, x_at={2: 1 / RT2, 3: 1 / RT2})


def epipersquare2():  # :1078-1092
    m = Model([0, 0, -1], [[1, 0, 0], [0, 1, 0]], [1 / RT2 / 2, 1 / RT2], -_I(3), np.zeros(3),
              [cs.EpiPerSquare(3)], obj_offset=-1.0)
    return m, dict(status="Optimal", primal_obj=-1 / RT2 - 1, x_at={1: 1 / RT2})


def epipersquare3():  # :1094-1106
    m = Model([0, 1, -1, -1], [[1, 0, 0, 0]], [0], -_I(4), np.zeros(4), [cs.EpiPerSquare(4)])
    return m, dict(status="Optimal", primal_obj=0.0, x_norm=0.0)


def epipersquare4():  # :1108-1141
    rows = [1, 1, 2, 2, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11]
    cols = [1, 7, 2, 3, 4, 2, 3, 5, 4, 7, 6, 5, 6, 7]
    vals = [1, -0.5, 1, 1, 1, -1, -1, -1, -1, -0.5, -1, -1, -1, -1]
    G = np.zeros((11, 7))
    for r, c_, v in zip(rows, cols, vals):
        G[r - 1, c_ - 1] = v
    c = np.zeros(7)
    c[0] = -1
    h = np.zeros(11)
    h[1] = 3
    cones = [cs.Nonnegative(2), cs.EpiPerSquare(3), cs.EpiPerSquare(3), cs.EpiPerSquare(3)]
    m = Model(c, np.zeros((0, 7)), [], G, h, cones)
    inv3 = 1 / 3
    r3 = inv3 * RT2
    return m, dict(
        status="Optimal", primal_obj=-1.0, x=[1, 1, 1, 1, RT2, RT2, 2],
        s=[0, 0, 1, 1, RT2, 1, 1, RT2, RT2, RT2, 2],
        z=[1, inv3, inv3, inv3, -r3, inv3, inv3, -r3, r3, r3, -2 * inv3],
    )


def _wsos_box(l, u, d):
    return pu.interpolate(pu.BoxDomain(l, u), d)


def wsosinterpnonnegative1():  # :2477-2494
    U, pts, Ps = _wsos_box([0, 0], [1, 1], 2)
    x, y = pts[:, 0], pts[:, 1]
    h = x**4 + x**2 * y**2 + 4 * y**2 + 4
    m = Model([-1], np.zeros((0, 1)), [], np.ones((U, 1)), h, [cs.WSOSInterpNonnegative(U, Ps)])
    return m, dict(status="Optimal", primal_obj=-4.0, x=[4.0])


def wsosinterpnonnegative2():  # :2496-2513
    U, pts, Ps = _wsos_box([0, 0], [3, 3], 2)
    x, y = pts[:, 0], pts[:, 1]
    h = (x - 2) ** 2 + (x * y - 3) ** 2
    m = Model([-1], np.zeros((0, 1)), [], np.ones((U, 1)), h, [cs.WSOSInterpNonnegative(U, Ps)])
    return m, dict(status="Optimal", primal_obj=0.0, x=[0.0])


def wsosinterpnonnegative3():  # :2515-2534
    U, pts, Ps = _wsos_box([0, 0], [3, 3], 2)
    x, y = pts[:, 0], pts[:, 1]
    c = (x - 2) ** 2 + (x * y - 3) ** 2
    m = Model(c, np.ones((1, U)), [1], -_I(U), np.zeros(U), [cs.WSOSInterpNonnegative(U, Ps, use_dual=True)])
    return m, dict(status="Optimal", primal_obj=0.0)


def wsosinterppossemideftri1():  # :2574-2592  H = d2/dx2 (x+1)^2 (x-1)^2 = 12x^2 - 4
    U, pts, Ps = pu.interpolate(pu.BoxDomain([-1.0], [1.0]), 1)
    x = pts[:, 0]
    h = 12 * x**2 - 4
    m = Model([-1], np.zeros((0, 1)), [], np.ones((U, 1)), h, [cs.wsos_psd(1, U, Ps)])
    return m, dict(status="Optimal", primal_obj=4.0, x=[-4.0])


def wsosinterppossemideftri2():  # :2594-2613  Hessian of x1^4 - 3 x2^2
    U, pts, Ps = pu.interpolate(pu.FreeDomain(2), 1)
    x1 = pts[:, 0]
    H = {(1, 1): 12 * x1**2, (2, 1): np.zeros(U), (2, 2): -6 * np.ones(U)}
    h = np.concatenate([H[(i, j)] for i in (1, 2) for j in range(1, i + 1)])
    cs.scale_svec_incr(h, RT2, U)
    G = np.vstack([np.ones((U, 1)), np.zeros((U, 1)), np.ones((U, 1))])
    m = Model([-1], np.zeros((0, 1)), [], G, h, [cs.wsos_psd(2, U, Ps)])
    return m, dict(status="Optimal", primal_obj=6.0, x=[-6.0])


def wsosinterppossemideftri3():  # :2615-2638
    U, pts, Ps = pu.interpolate(pu.FreeDomain(1), 3)
    x = pts[:, 0]
    M = [[x + 2 * x**3, np.ones(U)], [-(x**2) + 2, 3 * x**2 - x + 1]]
    MtM = {}
    for i in range(2):
        for j in range(2):
            MtM[(i, j)] = (M[0][i] * M[0][j] + M[1][i] * M[1][j]) / 10
    h = np.concatenate([MtM[(i, j)] for i in range(2) for j in range(i + 1)])
    cs.scale_svec_incr(h, RT2, U)
    m = Model([], np.zeros((0, 0)), [], np.zeros((h.size, 0)), h, [cs.wsos_psd(2, U, Ps)])
    return m, dict(status="Optimal", primal_obj=0.0)


def doublynonnegativetri1():  # test/nativeinstances.jl:558-576 (the loop body pins use_dual = false)
    m = Model([0, 1, 0], [[1, 0, 0], [0, 0, 1]], [1, 1], -_I(3), np.zeros(3), [cs.DoublyNonnegativeTri(3)])
    return m, dict(status="Optimal", primal_obj=0.0, x=[1, 0, 1], s=[1, 0, 1])


def doublynonnegativetri2():  # :578-591
    m = Model([0, -1, 0], [[1, 0, 0], [0, 0, 1]], [1.0, 1.5], -_I(3), [-0.5, 0, -0.5], [cs.DoublyNonnegativeTri(3)])
    return m, dict(status="Optimal", primal_obj=-1.0, x_at={1: 1.0})


def possemideftrisparse1():  # test/nativeinstances.jl:608-628 (dense implementation, real)
    cone = cs.psd_sparse(2, [1, 2, 2], [1, 1, 2])
    m = Model([0, -1, 0], [[1, 0, 0], [0, 0, 1]], [0.5, 1], -_I(3), np.zeros(3), [cone])
    return m, dict(status="Optimal", primal_obj=-1.0, x_at={1: 1.0})


def possemideftrisparse4():  # :703-733
    G = np.zeros((10, 1))
    G[[0, 1, 2, 5, 9], 0] = -1
    h = np.zeros(10)
    h[[3, 4, 6, 7, 8]] = RT2 * np.array([1, 1, 1, -1, 1.0])
    cone = cs.psd_sparse(5, [1, 2, 3, 4, 4, 4, 5, 5, 5, 5], [1, 2, 3, 1, 2, 4, 1, 2, 3, 5])
    m = Model([1], np.zeros((0, 1)), [], G, h, [cone])
    inv6, invrt6 = 1 / 6, 1 / (RT2 * RT3)
    s = [RT3, RT3, RT3, RT2, RT2, RT3, RT2, -RT2, RT2, RT3]
    z = [inv6, inv6, inv6, 0, 0, 0, -invrt6, invrt6, -invrt6, 0.5]
    return m, dict(status="Optimal", primal_obj=RT3, s=s, z=z)


def possemideftrisparse5():  # :735-784 (use_dual = true)
    G = np.vstack([-_I(6), np.zeros((6, 6))])
    G[0, 1:] = 0.5
    h = np.concatenate([np.zeros(6), RT2 * np.array([1, 1, 0, 1, -1, 1.0])])
    cone = cs.psd_sparse(5, [1, 2, 3, 4, 5, 4, 4, 4, 5, 5, 5], [1, 2, 3, 4, 5, 1, 2, 3, 1, 2, 3], use_dual=True)
    c = np.concatenate([[1.0], np.zeros(5)])
    m = Model(c, np.zeros((0, 6)), [], G, h, [cs.Nonnegative(1), cone])
    invrt2, invrt3 = 1 / RT2, 1 / RT3
    invrt6 = invrt2 * invrt3
    s = [0, invrt2 + invrt3, invrt2 + invrt3, invrt3, RT2, RT3, RT2, RT2, 0, RT2, -RT2, RT2]
    z = [1, 0.5, 0.5, 0.5, 0.5, 0.5, -0.5, -0.5, 0, -invrt6, invrt6, -invrt6]
    return m, dict(status="Optimal", primal_obj=RT2 + RT3, s=s, z=z)


def wsosinterpepinormeucl1():  # test/nativeinstances.jl:2715-2737: min t : t^2 >= x^4 on [-1, 1], t constant
    U, pts, Ps = _wsos_box([-1], [1], 1)
    assert U == 3
    h = np.concatenate([np.zeros(U), pts[:, 0] ** 2])
    G = np.vstack([-_I(U), np.zeros((U, U))])
    m = Model(np.ones(U), [[1, -1, 0], [1, 0, -1]], [0, 0], G, h, [cs.wsos_epinormeucl(2, U, Ps)])
    return m, dict(status="Optimal", primal_obj=float(U), x=np.ones(U))


def wsosinterpepinormeucl2():  # :2739-2766: min t : t^2 >= x^4 + (x - 1)^2 on [-1, 1]
    U, pts, Ps = _wsos_box([-1], [1], 1)
    x = pts[:, 0]
    h = np.concatenate([np.zeros(U), x**2, x - 1])
    G = np.vstack([-_I(U), np.zeros((U, U)), np.zeros((U, U))])
    m = Model(np.ones(U), [[1, -1, 0], [1, 0, -1]], [0, 0], G, h, [cs.wsos_epinormeucl(3, U, Ps)])
    return m, dict(status="Optimal", primal_obj=math.sqrt(5.0) * U, x=np.full(U, math.sqrt(5.0)))


def wsosinterpepinormone1():  # test/nativeinstances.jl:2640-2660: min t : t >= |x^2| on [-1, 1], t constant
    U, pts, Ps = _wsos_box([-1], [1], 1)
    h = np.concatenate([np.zeros(U), pts[:, 0] ** 2])
    G = np.vstack([-_I(U), np.zeros((U, U))])
    m = Model(np.ones(U), [[1, -1, 0], [1, 0, -1]], [0, 0], G, h, [cs.wsos_epinormone(2, U, Ps)])
    return m, dict(status="Optimal", primal_obj=float(U), x=np.ones(U))


def wsosinterpepinormone2():  # :2662-2688: min t : t >= |x^2| + |x - 1| on [-1, 1]
    U, pts, Ps = _wsos_box([-1], [1], 1)
    x = pts[:, 0]
    h = np.concatenate([np.zeros(U), x**2, x - 1])
    G = np.vstack([-_I(U), np.zeros((U, U)), np.zeros((U, U))])
    m = Model(np.ones(U), [[1, -1, 0], [1, 0, -1]], [0, 0], G, h, [cs.wsos_epinormone(3, U, Ps)])
    return m, dict(status="Optimal", primal_obj=3.0 * U, x=np.full(U, 3.0))


KNOWN_ANSWER = {
    f.__name__: f
    for f in (
        dimension1, primalinfeas1, primalinfeas2, dualinfeas2, dualinfeas3, nonnegative4,
        possemideftri1, possemideftri2, possemideftri8, possemideftri9,
        epinorminf1, epinorminf2, epinorminf3, epinorminf3_dual, epinorminf4,
        doublynonnegativetri1, doublynonnegativetri2,
        possemideftrisparse1, possemideftrisparse4, possemideftrisparse5,
        epinormeucl1, epinormeucl2, epinormeucl3,
        epipersquare1, epipersquare2, epipersquare3, epipersquare4,
        wsosinterpnonnegative1, wsosinterpnonnegative2, wsosinterpnonnegative3,
        wsosinterppossemideftri1, wsosinterppossemideftri2, wsosinterppossemideftri3,
        wsosinterpepinormeucl1, wsosinterpepinormeucl2,
        wsosinterpepinormone1, wsosinterpepinormone2,
    )
}


# ---------------------------------------------------------------------------
# BASELINE.json configs (synthetic)
# ---------------------------------------------------------------------------
def linearopt(m: int = 500, n: int = 1000, seed: int = 1) -> Model:
    """cfg1 — examples/linearopt/native.jl:22-37 (dense, nz_frac = 1)."""
    rng = np.random.default_rng(seed)
    A = 10.0 * rng.random((m, n))
    b = A.sum(axis=1)
    c = rng.random(n)
    return Model(c, A, b, -_I(n), np.zeros(n), [cs.Nonnegative(n)])


def nearestpsd(side: int = 300, seed: int = 1) -> Model:
    """cfg2 variant A (use_completable = false) — examples/nearestpsd/JuMP.jl:34-89
    lowered the way MOI does (svec with sqrt2 scaling; one equality row)."""
    rng = np.random.default_rng(seed)
    sparsity = min(3.0 / side, 1.0)
    mask = np.tril(rng.random((side, side)) < sparsity, -1)
    rows, cols = np.nonzero(mask)
    # tril(sprandn) + Diagonal(randn): strictly-lower pattern entries plus every diagonal entry
    rows = np.concatenate([rows, np.arange(side)])
    cols = np.concatenate([cols, np.arange(side)])
    order = np.lexsort((rows, cols))  # column-major like findnz
    rows, cols = rows[order], cols[order]
    vals = rng.standard_normal(rows.size)
    is_diag = rows == cols
    nvar = rows.size
    # max 2<A,X> - sum_diag A_kk X_kk   ->   min -(...)
    c = -(2 * vals - np.where(is_diag, vals, 0.0))
    A = is_diag.astype(np.float64)[None, :]
    b = [1.0]
    dim = cs.svec_length(side)
    G = np.zeros((dim, nvar))
    for k in range(nvar):
        i, j = rows[k] + 1, cols[k] + 1  # i >= j
        idx = cs.svec_idx(i, j) - 1
        G[idx, k] = -1.0 if i == j else -RT2
    return Model(c, A, b, G, np.zeros(dim), [cs.PosSemidefTri(dim)])


def nearestdnn(side: int = 12, seed: int = 1) -> Model:
    """Doubly nonnegative relaxation test model (synthetic): max <C, X> s.t. tr(X) = 1, <J, X> = side / 2, X in the
    doubly nonnegative cone, lowered like nearestpsd (svec with sqrt2 scaling)."""
    rng = np.random.default_rng(seed)
    dim = cs.svec_length(side)
    Cm = rng.standard_normal((side, side))
    Cm = 0.5 * (Cm + Cm.T)
    c = -cs.smat_to_svec(Cm)
    # variables are the svec coordinates themselves: G = -I, h = 0
    tr = cs.smat_to_svec(np.eye(side))
    ones = cs.smat_to_svec(np.ones((side, side)))
    A = np.vstack([tr, ones])
    b = [1.0, 1.0 + 0.25 * (side - 1)]
    return Model(c, A, b, -_I(dim), np.zeros(dim), [cs.DoublyNonnegativeTri(dim)])


def polynormeucl(n: int = 1, halfdeg: int = 3, R: int = 3, seed: int = 1) -> Model:
    """Seeded model over WSOSInterpEpiNormEucl in the shape of wsosinterpepinormeucl2: the smallest constant t with
    t^2 >= sum_i f_i(x)^2 on the box [-1, 1]^n for R - 1 random polynomials given by their values at the points."""
    rng = np.random.default_rng(seed)
    U, pts, Ps = pu.interpolate(pu.BoxDomain(-np.ones(n), np.ones(n)), halfdeg, sample=False)
    fs = [rng.standard_normal(n + 1) for _ in range(R - 1)]
    vals = [f[0] + pts @ f[1:] for f in fs]  # affine polynomials: inside the degree the basis represents
    h = np.concatenate([np.zeros(U)] + vals)
    G = np.vstack([-np.ones((U, 1))] + [np.zeros((U, 1))] * (R - 1))
    return Model([1.0], np.zeros((0, 1)), [], G, h, [cs.wsos_epinormeucl(R, U, Ps)])


def polynormone(n: int = 1, halfdeg: int = 3, R: int = 3, seed: int = 1) -> Model:
    """polynormeucl with the 1-norm: the smallest constant t with t >= sum_i |f_i(x)| on the box (WSOSInterpEpiNormOne)."""
    rng = np.random.default_rng(seed)
    U, pts, Ps = pu.interpolate(pu.BoxDomain(-np.ones(n), np.ones(n)), halfdeg, sample=False)
    fs = [rng.standard_normal(n + 1) for _ in range(R - 1)]
    vals = [f[0] + pts @ f[1:] for f in fs]
    h = np.concatenate([np.zeros(U)] + vals)
    G = np.vstack([-np.ones((U, 1))] + [np.zeros((U, 1))] * (R - 1))
    return Model([1.0], np.zeros((0, 1)), [], G, h, [cs.wsos_epinormone(R, U, Ps)])


def sparsepsd_maxeig(side: int = 20, seed: int = 1) -> Model:
    """The model of test/nativeinstances.jl:657-701 (possemideftrisparse3, real case) on a seeded pattern: the largest
    eigenvalue of a sparse symmetric matrix as a conic program over the sparse PSD cone with the dual barrier."""
    rng = np.random.default_rng(seed)
    mask = np.tril(rng.random((side, side)) < 1.0 / math.sqrt(side), -1) | np.eye(side, dtype=bool)
    cols, rows = np.nonzero(mask.T)
    vals = rng.random(rows.size) + (rows == cols)
    dim = rows.size
    c = -np.where(rows == cols, vals, RT2 * vals)
    A = (rows == cols).astype(np.float64)[None, :]
    cone = cs.psd_sparse(side, rows + 1, cols + 1, use_dual=True)
    return Model(c, A, [1.0], -_I(dim), np.zeros(dim), [cone])


def polymin(n: int = 2, halfdeg: int = 30, seed: int = 1) -> Model:
    """cfg3 — examples/polymin/native.jl:48-113 with use_primal = false, use_wsos = true
    (dual form: c = interp values, A = 1', b = 1, G = -I, WSOS cone with use_dual = true)."""
    rng = np.random.default_rng(seed)
    U, pts, Ps = pu.interpolate(pu.BoxDomain(-np.ones(n), np.ones(n)), halfdeg)
    c = rng.standard_normal(U)
    return Model(c, np.ones((1, U)), [1.0], -_I(U), np.zeros(U), [cs.WSOSInterpNonnegative(U, Ps, use_dual=True)])


def portfolio_socs(K: int = 100_000, d: int = 4, n: int = 2048, seed: int = 1) -> Model:
    """cfg4 — K scenario risk constraints ||S_j x||_2 <= gamma_j (EpiNormEucl(d)),
    1'x = 1, x >= 0 (the reference's portfolio has one big cone)."""
    rng = np.random.default_rng(seed)
    returns = rng.random(n)
    x0 = rng.random(n)
    x0 /= x0.sum()
    S = rng.standard_normal((K, d - 1, n))
    gammas = np.linalg.norm(S @ x0, axis=1) * 1.5 + 1e-3
    q = n + K * d
    G = np.zeros((q, n))
    h = np.zeros(q)
    G[:n, :] = -_I(n)
    cones = [cs.Nonnegative(n)]
    for j in range(K):
        r0 = n + j * d
        h[r0] = gammas[j]
        G[r0 + 1 : r0 + d, :] = -S[j]
        cones.append(cs.EpiNormEucl(d))
    return Model(-returns, np.ones((1, n)), [1.0], G, h, cones)


def portfolio_norminf(K: int = 100_000, d: int = 4, n: int = 2048, seed: int = 1) -> Model:
    """cfg4, EpiNormInf variant (BASELINE config 4 names "EpiNormEucl/EpiNormInf"): K scenario risk
    constraints ||S_j x||_inf <= gamma_j (EpiNormInf(d), the cone without sqrt oracles: every iteration
    takes the gemm branch of the Schur assembly, qrchol.jl:248-254), 1'x = 1, x >= 0."""
    rng = np.random.default_rng(seed)
    returns = rng.random(n)
    x0 = rng.random(n)
    x0 /= x0.sum()
    S = rng.standard_normal((K, d - 1, n))
    gammas = np.max(np.abs(S @ x0), axis=1) * 1.5 + 1e-3
    q = n + K * d
    G = np.zeros((q, n))
    h = np.zeros(q)
    G[:n, :] = -_I(n)
    cones = [cs.Nonnegative(n)]
    for j in range(K):
        r0 = n + j * d
        h[r0] = gammas[j]
        G[r0 + 1 : r0 + d, :] = -S[j]
        cones.append(cs.EpiNormInf(d))
    return Model(-returns, np.ones((1, n)), [1.0], G, h, cones)


def block_multicone(N: int = 8192, seed: int = 1, soc_dim: int = 8, psd_side: int = 16) -> Model:
    """cfg5 — p = 0, q = 2N split in quarters: Nonnegative / EpiNormEucl(8) /
    EpiPerSquare(8) / PosSemidefTri(side 16); G ~ N(0,1)/sqrt(N); strictly feasible
    primal-dual pair by construction so the status is Optimal."""
    rng = np.random.default_rng(seed)
    quarter = N // 2
    cones = [cs.Nonnegative(quarter)]
    cones += [cs.EpiNormEucl(soc_dim) for _ in range(quarter // soc_dim)]
    cones += [cs.EpiPerSquare(soc_dim) for _ in range(quarter // soc_dim)]
    psd_dim = cs.svec_length(psd_side)
    n_psd = max(1, quarter // psd_dim)
    cones += [cs.PosSemidefTri(psd_dim) for _ in range(n_psd)]
    q = sum(c.dim for c in cones)
    G = rng.standard_normal((q, N)) / math.sqrt(N)
    x0 = rng.standard_normal(N)
    s0 = np.concatenate([c.initial_point() for c in cones])
    z0 = s0.copy()  # initial points are central: z0 = -g(s0) = s0
    # perturb inside the cones (small relative noise keeps strict feasibility)
    s0 = s0 + 0.1 * _interior_noise(cones, rng)
    z0 = z0 + 0.1 * _interior_noise(cones, rng)
    h = G @ x0 + s0
    c = -G.T @ z0
    return Model(c, np.zeros((0, N)), [], G, h, cones)


def _interior_noise(cones, rng) -> np.ndarray:
    out = []
    for c in cones:
        if c.type_code == cs.NONNEGATIVE:
            out.append(rng.random(c.dim))
        elif c.type_code == cs.POSSEMIDEFTRI:
            B = rng.standard_normal((c.side, c.side)) / math.sqrt(c.side)
            out.append(cs.smat_to_svec(B @ B.T))
        else:  # SOC / RSOC: noise on w only, bounded well inside
            v = np.zeros(c.dim)
            k0 = 1 if c.type_code == cs.EPINORMEUCL else 2
            w = rng.standard_normal(c.dim - k0)
            v[k0:] = 0.5 * w / max(np.linalg.norm(w), 1.0)
            out.append(v)
    return np.concatenate(out)
