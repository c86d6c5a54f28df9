"""Design prototype (numpy/scipy, test infrastructure only) of the smoothed-aggregation multigrid
preconditioner that replaces the reference's `pc_type gamg` on the Poisson pseudo-timestep
(movement/solver_parameters.py:68-71, movement/monge_ampere.py:441-448).

It mirrors, step by step, the data-parallel algorithm of movement_b200/csrc/mv_amg.cu:
  * strength graph  |a_ij| > theta * sqrt(a_ii a_jj)
  * distance-2 maximal independent set by Luby rounds with a hashed priority  -> roots
  * aggregates = root + strong neighbours; leftovers join their strongest aggregated neighbour
  * P = (I - omega D^-1 A) T,  T = piecewise constant,  omega = 4/(3 rho)
  * Galerkin A_c = P^T A P
  * V(1,1)-cycle with a Chebyshev(k) smoother on D^-1 A, dense (shifted) inverse on the coarsest
It is used to choose parameters and as the checker of the device hierarchy in tests/.
"""
from __future__ import annotations

import numpy as np
import scipy.sparse as sp


def hash32(i):
    x = (np.asarray(i, dtype=np.uint64) + np.uint64(0x9E3779B9)) & np.uint64(0xFFFFFFFF)
    for mul in (0x85EBCA6B, 0xC2B2AE35):
        x ^= x >> np.uint64(16)
        x = (x * np.uint64(mul)) & np.uint64(0xFFFFFFFF)
    x ^= x >> np.uint64(16)
    return x


def strength_graph(A, theta):
    A = A.tocsr()
    d = A.diagonal()
    C = A.tocoo()
    keep = (C.row != C.col) & (np.abs(C.data) > theta * np.sqrt(np.abs(d[C.row] * d[C.col])))
    S = sp.csr_matrix((np.abs(C.data[keep]), (C.row[keep], C.col[keep])), shape=A.shape)
    return S


def mis2_aggregate(S, active=None):
    """Luby MIS-2 on the strength graph, then two attachment passes. Returns agg [n] (-1 = none)."""
    n = S.shape[0]
    S = S.tocsr()
    if active is None:
        active = np.ones(n, dtype=bool)
    # priority: (hash, index) packed so that ties are impossible; 0 = not a candidate
    prio = ((hash32(np.arange(n)) << np.uint64(32)) | np.arange(1, n + 1, dtype=np.uint64))
    UND, IN, OUT = 0, 1, 2
    state = np.where(active, UND, OUT).astype(np.int8)
    indptr, indices = S.indptr, S.indices
    rows = np.repeat(np.arange(n), np.diff(indptr))

    def nbr_max(v):
        out = v.copy()
        np.maximum.at(out, rows, v[indices])
        return out

    rounds = 0
    while (state == UND).any():
        rounds += 1
        p = np.where(state == UND, prio, np.uint64(0))
        m2 = nbr_max(nbr_max(p))
        newin = (state == UND) & (m2 == prio)
        state[newin] = IN
        f = (state == IN).astype(np.uint64)
        near = nbr_max(nbr_max(f)) > 0
        state[(state == UND) & near] = OUT
    roots = np.where(state == IN)[0]
    agg = np.full(n, -1, dtype=np.int64)
    agg[roots] = np.arange(len(roots))
    # pass 1: neighbours of a root (at most one root is adjacent to any vertex)
    rootid = np.where(state == IN, agg, -1)
    nb = np.full(n, -1, dtype=np.int64)
    np.maximum.at(nb, rows, rootid[indices])
    take = (agg < 0) & active & (nb >= 0)
    agg[take] = nb[take]
    # pass 2: leftovers join the aggregate of their strongest already-aggregated neighbour
    agg1 = agg.copy()
    left = np.where((agg < 0) & active)[0]
    for i in left:
        js = indices[indptr[i]:indptr[i + 1]]
        w = S.data[indptr[i]:indptr[i + 1]]
        ok = agg1[js] >= 0
        if ok.any():
            k = np.lexsort((-js[ok], w[ok]))[-1]      # strongest, ties -> lowest index
            agg[i] = agg1[js[ok][k]]
    # isolated vertices (no strong neighbour at all): singleton aggregates
    lone = np.where((agg < 0) & active)[0]
    agg[lone] = len(roots) + np.arange(len(lone))
    return agg, int(agg.max()) + 1, rounds


def rho_DinvA(A, iters=15):
    d = A.diagonal()
    n = A.shape[0]
    x = (hash32(np.arange(n)).astype(np.float64) / 2 ** 32) - 0.5
    lam = 1.0
    for _ in range(iters):
        y = (A @ x) / d
        lam = np.linalg.norm(y) / np.linalg.norm(x)
        x = y / np.linalg.norm(y)
    return lam


class Level:
    pass


def build_hierarchy(A, theta=0.08, max_coarse=200, max_levels=10, smooth_P=True, fixed=None, verbose=False):
    levels = []
    A = A.tocsr()
    while True:
        L = Level()
        L.A = A
        L.d = A.diagonal()
        L.rho = rho_DinvA(A) * 1.1
        levels.append(L)
        n = A.shape[0]
        if n <= max_coarse or len(levels) >= max_levels:
            break
        S = strength_graph(A, theta)
        active = None
        if fixed is not None and len(levels) == 1:
            active = ~fixed
        agg, nc, rounds = mis2_aggregate(S, active)
        sel = agg >= 0
        T = sp.csr_matrix((np.ones(sel.sum()), (np.where(sel)[0], agg[sel])), shape=(n, nc))
        if smooth_P:
            omega = 4.0 / (3.0 * L.rho)
            P = T - omega * sp.diags(1.0 / L.d) @ (A @ T)
            if active is not None:
                P = sp.diags(active.astype(float)) @ P
        else:
            P = T
        L.P = P.tocsr()
        L.agg = agg
        Ac = (L.P.T @ A @ L.P).tocsr()
        Ac.eliminate_zeros()
        if verbose:
            print(f"  level {len(levels)-1}: n={n} nnz={A.nnz} -> nc={nc} nnz_c={Ac.nnz} ({Ac.nnz/nc:.1f}/row) "
                  f"mis rounds {rounds} rho={L.rho:.3f}")
        A = Ac
    return levels


def cheby(L, b, x, degree, ratio):
    """x <- x + p(D^-1 A) D^-1 (b - A x): Chebyshev on [rho/ratio, rho]; x=None means zero guess."""
    lmax, lmin = L.rho, L.rho / ratio
    theta, delta = 0.5 * (lmax + lmin), 0.5 * (lmax - lmin)
    sigma = theta / delta
    rho_k = 1.0 / sigma
    r = b.copy() if x is None else b - L.A @ x
    dk = (r / L.d) / theta
    x = dk.copy() if x is None else x + dk
    for _ in range(degree - 1):
        r = r - L.A @ dk
        rho_n = 1.0 / (2.0 * sigma - rho_k)
        dk = rho_n * rho_k * dk + (2.0 * rho_n / delta) * (r / L.d)
        rho_k = rho_n
        x = x + dk
    return x


def vcycle(levels, l, b, degree=2, ratio=4.0, singular=True):
    L = levels[l]
    if l == len(levels) - 1:
        if not hasattr(L, "inv"):
            n = L.A.shape[0]
            D = L.A.toarray()
            if singular:
                D = D + np.full((n, n), np.abs(L.d).mean() / n)
            L.inv = np.linalg.inv(D)
        return L.inv @ b
    x = cheby(L, b, None, degree, ratio)
    r = b - L.A @ x
    e = vcycle(levels, l + 1, L.P.T @ r, degree, ratio, singular)
    x = x + L.P @ e
    return cheby(L, b, x, degree, ratio)


def pcg(A, b, M, rtol=1e-12, maxit=500, project=True):
    x = np.zeros_like(b)
    if project:
        b = b - b.mean()
    r = b.copy()
    z = M(r)
    p = z.copy()
    rz = r @ z
    ref = rz
    it = 0
    hist = [1.0]
    while it < maxit and rz > rtol * rtol * ref:
        q = A @ p
        al = rz / (p @ q)
        x += al * p
        r -= al * q
        if project:
            r -= r.mean()
        z = M(r)
        rzn = r @ z
        p = z + (rzn / rz) * p
        rz = rzn
        it += 1
        hist.append(np.sqrt(rz / ref))
    return x, it, hist


if __name__ == "__main__":
    import sys
    import time

    import movement_oracle as mo

    n = int(sys.argv[1]) if len(sys.argv) > 1 else 24
    dim = int(sys.argv[2]) if len(sys.argv) > 2 else 3
    mesh = mo.unit_cube_mesh(n) if dim == 3 else mo.unit_square_mesh(n)
    vol, grads = mo.cell_geometry(mesh.coords, mesh.cells)
    K = mo.assemble_stiffness(mesh.num_vertices, mesh.cells, vol, grads)
    rng = np.random.default_rng(0)
    b = rng.standard_normal(K.shape[0])
    b -= b.mean()
    dinv = 1.0 / K.diagonal()
    _, itj, _ = pcg(K, b, lambda r: dinv * r, maxit=20000)
    print(f"n={n} dim={dim} V={K.shape[0]}: Jacobi-PCG {itj} its")
    for theta in (0.0, 0.08):
        for smooth in (True, False):
            t0 = time.time()
            lv = build_hierarchy(K, theta=theta, smooth_P=smooth, verbose=True)
            cx = sum(L.A.nnz for L in lv) / lv[0].A.nnz
            for degree, ratio in ((1, 4.0), (2, 4.0), (3, 8.0)):
                _, it, h = pcg(K, b, lambda r: vcycle(lv, 0, r, degree, ratio))
                print(f"theta={theta} smoothP={smooth} cheby({degree},{ratio}): {it} its, op-complexity {cx:.2f}, "
                      f"levels {len(lv)}, fine-spmv-equiv/it ~{(4*degree-1+1)*cx:.1f}  ({time.time()-t0:.1f}s)")
