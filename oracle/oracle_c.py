"""ctypes wrapper of the C/OpenMP restatement (oracle/movement_oracle_c.c).

TEST / MEASUREMENT INFRASTRUCTURE ONLY: used by tests/test_oracle_c.py (checked against the numpy
oracle) and by bench.py's cpu_baseline / --impl reference legs.  Never imported by movement_b200.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "_build", "libmovement_oracle_c.so")

MON_CONST, MON_BALL, MON_BALL_RAW, MON_GAUSS_HESS, MON_SINATAN3 = 0, 1, 2, 3, 4
_lib = None


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            subprocess.check_call(["make", "-s", "-C", HERE])
        lib = C.CDLL(LIB)
        vp, i64, dbl = C.c_void_p, C.c_int64, C.c_double
        lib.orc_ma_create.restype = vp
        lib.orc_ma_create.argtypes = [C.c_int, i64, i64, vp, vp, i64, vp, vp, vp, vp, vp, C.c_int]
        lib.orc_ma_destroy.argtypes = [vp]
        lib.orc_ma_step.restype = C.c_int
        lib.orc_ma_step.argtypes = [vp, C.c_int, vp, dbl, dbl, C.c_int, vp, vp, vp]
        lib.orc_ma_get.argtypes = [vp, C.c_int, vp]
        lib.orc_ma_threads.argtypes = [vp]
        _lib = lib
    return _lib


def boundary_facet_geometry(coords, cells, bfacets):
    """Vectorised: adjacent cell, outward unit normal and measure of every boundary facet."""
    V, d = coords.shape
    d1 = d + 1
    faces = np.concatenate([np.delete(cells, loc, axis=1) for loc in range(d1)]).astype(np.int64)
    fcell = np.tile(np.arange(cells.shape[0]), d1)
    fopp = np.concatenate([cells[:, loc] for loc in range(d1)])
    # Match every boundary facet with the cell facet that has the same (sorted) vertices.  Only facets made of boundary
    # vertices can match, and the boundary vertices are renumbered 0..nb-1 first: the key ((v0*nb)+v1)*nb+v2 then needs
    # nb^d < 2^63 instead of V^d (V^3 overflowed int64 above 126^3 cells; nb = 6n^2+2 stays far below at any size a
    # host can hold).
    bf = bfacets.astype(np.int64)
    bverts = np.unique(bf)
    nb = int(bverts.size)
    assert float(nb) ** d < 9e18, "facet keys would overflow int64"
    renum = np.full(V, -1, dtype=np.int64)
    renum[bverts] = np.arange(nb)
    rf = renum[faces]
    cand = np.flatnonzero((rf >= 0).all(axis=1))

    def key(f):
        f = np.sort(f, axis=1)
        k = f[:, 0].copy()
        for a in range(1, d):
            k = k * nb + f[:, a]
        return k

    fk = key(rf[cand])
    order = np.argsort(fk, kind="stable")
    bk = key(renum[bf])
    at = np.searchsorted(fk[order], bk)
    assert at.max(initial=0) < fk.size and np.array_equal(fk[order][at], bk), "a boundary facet belongs to no cell"
    pos = cand[order[at]]
    cell, opp = fcell[pos], fopp[pos]
    fx = coords[bfacets]
    if d == 2:
        e = fx[:, 1] - fx[:, 0]
        area = np.linalg.norm(e, axis=1)
        n = np.stack([e[:, 1], -e[:, 0]], axis=1) / area[:, None]
    else:
        nn = np.cross(fx[:, 1] - fx[:, 0], fx[:, 2] - fx[:, 0])
        ln = np.linalg.norm(nn, axis=1)
        area = 0.5 * ln
        n = nn / ln[:, None]
    flip = np.einsum("fa,fa->f", n, fx[:, 0] - coords[opp]) < 0
    n[flip] *= -1
    return cell.astype(np.int32), n, area


def axis_aligned_bcmask(V, d, bfacets, btags, normal, area, fixed=()):
    """uint16 mask of the L2-projector Dirichlet dofs (movement/monge_ampere.py:246-272)."""
    mask = np.zeros(V, dtype=np.uint16)
    for t in np.unique(btags):
        sel = btags == t
        nodes = np.unique(bfacets[sel])
        if t in fixed:
            mask[nodes] |= (1 << d) - 1
            continue
        absn = (np.abs(normal[sel]) * area[sel, None]).sum(axis=0)
        iszero = np.isclose(absn, 0.0)
        if iszero.sum() != d - 1:
            raise NotImplementedError("non-axis-aligned boundary segment")
        mask[nodes] |= 1 << int(np.where(~iszero)[0][0])
    return mask


class CRelaxation:
    """movement/monge_ampere.py:488-540 on the CPU with OpenMP Jacobi-PCG inner solves."""

    def __init__(self, coords, cells, bfacets, btags, monitor_kind, monitor_params, pseudo_timestep=0.1,
                 inner_rtol=1e-12, nthreads=0, fixed=()):
        lib = load()
        self.coords = np.ascontiguousarray(coords, dtype=np.float64)
        self.cells = np.ascontiguousarray(cells, dtype=np.int32)
        bf = np.ascontiguousarray(bfacets, dtype=np.int32)
        V, d = self.coords.shape
        fcell, fn, fa = boundary_facet_geometry(self.coords, self.cells, bf)
        mask = axis_aligned_bcmask(V, d, bf, np.asarray(btags), fn, fa, fixed)
        fn = np.ascontiguousarray(fn); fa = np.ascontiguousarray(fa)
        self.V, self.dim, self.C = V, d, self.cells.shape[0]
        p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
        self.h = lib.orc_ma_create(d, V, self.C, p(self.coords), p(self.cells), bf.shape[0], p(bf), p(fcell), p(fn), p(fa),
                                   p(mask), int(nthreads))
        self.kind = int(monitor_kind)
        self.params = np.zeros(8)
        self.params[:len(monitor_params)] = monitor_params
        self.dt, self.inner_rtol = float(pseudo_timestep), float(inner_rtol)
        self.trace, self.cg_iterations = [], []
        self.threads = lib.orc_ma_threads(self.h)

    def __del__(self):
        if getattr(self, "h", None) and _lib is not None:
            _lib.orc_ma_destroy(self.h)
            self.h = None

    def step(self, update=True):
        tr = np.zeros(4)
        it = np.zeros(3, dtype=np.int32)
        nt = C.c_int64(0)
        load().orc_ma_step(self.h, self.kind, self.params.ctypes.data_as(C.c_void_p), self.dt, self.inner_rtol,
                           1 if update else 0, tr.ctypes.data_as(C.c_void_p), it.ctypes.data_as(C.c_void_p), C.byref(nt))
        return tr, it, nt.value

    def move(self, maxiter=1000, rtol=1e-8, dtol=2.0):
        i = 0
        for i in range(maxiter):
            tr, it, _ = self.step(update=False)
            self.trace.append((i, tr[0], tr[1], tr[2]))
            if i == 0:
                r0 = tr[2]
            if tr[2] < rtol:
                break
            if tr[2] > dtol * r0 or i == maxiter - 1:
                raise RuntimeError("not converged")
            tr, it, _ = self.step(update=True)  # re-evaluates the first half (cheap) and advances
            self.cg_iterations.append(tuple(it))
        return i

    def get(self, what):
        d, V = self.dim, self.V
        shape = {0: (V, d), 1: (V,), 2: (V, d, d), 3: (V, d), 4: (V,)}[what]
        out = np.empty(shape)
        load().orc_ma_get(self.h, what, out.ctypes.data_as(C.c_void_p))
        return out

    @property
    def x(self):
        return self.get(0)
