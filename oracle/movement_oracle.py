"""CPU oracle for the Movement mesh-relocation hot path (numpy / scipy, fp64).

TEST INFRASTRUCTURE ONLY.  Nothing under ``movement_b200/`` may import this module;
only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs use it, and only as the checker / the timed CPU baseline.

Parity status: the reference (mesh-adaptation/movement) delegates every arithmetic
operation to Firedrake / PETSc / UFL, none of which exist in this container, so the
reference itself cannot be executed here.  This file is a restatement of the
reference's algorithm from its sources (citations ``file:line`` are relative to
``/root/reference``) using the closed forms that hold for affine P1 simplices and
*exact* (direct, sparse-LU) inner solves.  It is pinned against every known answer the
reference's tests and demos hold for the path (tests/test_oracle_golden.py):
printed iteration-0 / final traces of the demos, the tangling KAT counts, the spring
edge KATs, the Laplacian and spring 11-line traces, the quasi-Newton iteration counts.
Parity at 1e-10 against a live Firedrake run is **unpinned in this container**.

Layouts follow Firedrake's ``dat.data`` arrays: scalar [V], vector [V,d], tensor
[V,d,d], P0 [C]; cell-node map [C,d+1] int32.
"""
from __future__ import annotations

import math

import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spla
from scipy.special import roots_jacobi

# --------------------------------------------------------------------------------------
# Mesh generators (Firedrake's UnitSquareMesh / UnitCubeMesh layouts)
# --------------------------------------------------------------------------------------


class OMesh:
    """Plain container: coords [V,d] f64, cells [C,d+1] i32, boundary facets + tags.

    periodic=True: the Firedrake layout of a periodic mesh (movement/mover.py:70-75, monge_ampere.py:236-239) --
    ``cells`` is the PERIODIC P1 cell-node map (vertices identified across the seam) while the coordinate field is
    DG1, one copy of every vertex per cell: coords [C*(d+1), d] with coordinate cell-node map
    ``ccells[c, j] = c*(d+1) + j``; there are no boundary facets."""

    def __init__(self, coords, cells, bfacets, btags, periodic=False):
        self.coords = np.ascontiguousarray(coords, dtype=np.float64)
        if self.coords.ndim == 1:
            self.coords = self.coords[:, None]
        self.cells = np.ascontiguousarray(cells, dtype=np.int32)
        self.dim = self.coords.shape[1]
        self.bfacets = np.ascontiguousarray(bfacets, dtype=np.int32).reshape(-1, self.dim)
        self.btags = np.ascontiguousarray(btags, dtype=np.int32)
        self.periodic = bool(periodic)
        if self.periodic:
            C, d1 = self.cells.shape
            assert self.coords.shape[0] == C * d1
            self.ccells = np.arange(C * d1, dtype=np.int32).reshape(C, d1)
            # P1 vertex -> coordinate node: the copy held by the LAST cell that contains the vertex (Firedrake's
            # interpolation of a coordinate expression into the periodic P1 space writes cell by cell, last one wins)
            self.rep = np.zeros(int(self.cells.max()) + 1, dtype=np.int64)
            self.rep[self.cells.ravel()] = np.arange(C * d1)
        else:
            self.ccells = self.cells
            self.rep = None

    @property
    def num_vertices(self):
        return int(self.cells.max()) + 1 if self.periodic else self.coords.shape[0]

    def to_coordinate_nodes(self, g):
        """P1 nodal field -> values at the coordinate nodes (identity on a P1 mesh; CG -> DG interpolation of
        movement/monge_ampere.py:236-239 on a periodic one)."""
        return g[self.cells].reshape((-1,) + g.shape[1:]) if self.periodic else g

    def at_vertices(self, x):
        """coordinate field -> one position per P1 vertex (for nodal monitor evaluation)."""
        return x[self.rep] if self.periodic else x

    @property
    def num_cells(self):
        return self.cells.shape[0]


def unit_square_mesh(nx, ny=None):
    """'left'-diagonal UnitSquareMesh: vertex (i,j) -> i*(ny+1)+j at (xs[i], ys[j]);
    quad corners q0=(i,j) q1=(i,j+1) q2=(i+1,j+1) q3=(i+1,j); triangles [q0,q1,q3],
    [q1,q2,q3]; tags 1: x=0, 2: x=1, 3: y=0, 4: y=1."""
    ny = nx if ny is None else ny
    xs = np.linspace(0.0, 1.0, nx + 1)
    ys = np.linspace(0.0, 1.0, ny + 1)
    X, Y = np.meshgrid(xs, ys, indexing="ij")
    coords = np.stack([X.ravel(), Y.ravel()], axis=1)
    ii, jj = np.meshgrid(np.arange(nx), np.arange(ny), indexing="ij")
    ii = ii.ravel()
    jj = jj.ravel()
    q0 = ii * (ny + 1) + jj
    q1 = q0 + 1
    q3 = q0 + (ny + 1)
    q2 = q3 + 1
    cells = np.empty((2 * nx * ny, 3), dtype=np.int32)
    cells[0::2] = np.stack([q0, q1, q3], axis=1)
    cells[1::2] = np.stack([q1, q2, q3], axis=1)
    bf, bt = [], []
    j = np.arange(ny)
    bf.append(np.stack([j, j + 1], axis=1)); bt.append(np.full(ny, 1))                       # x = 0
    bf.append(np.stack([nx * (ny + 1) + j, nx * (ny + 1) + j + 1], axis=1)); bt.append(np.full(ny, 2))
    i = np.arange(nx)
    bf.append(np.stack([i * (ny + 1), (i + 1) * (ny + 1)], axis=1)); bt.append(np.full(nx, 3))  # y = 0
    bf.append(np.stack([i * (ny + 1) + ny, (i + 1) * (ny + 1) + ny], axis=1)); bt.append(np.full(nx, 4))
    return OMesh(coords, cells, np.concatenate(bf), np.concatenate(bt))


_KUHN = np.array(
    [[0, 1, 3, 7], [0, 1, 7, 5], [0, 5, 7, 4], [0, 3, 2, 7], [0, 6, 4, 7], [0, 2, 6, 7]]
)


def unit_cube_mesh(nx, ny=None, nz=None):
    """UnitCubeMesh, x fastest: v0=k(ny+1)(nx+1)+j(nx+1)+i; six Kuhn tets per hex;
    tags 1/2: x=0/1, 3/4: y=0/1, 5/6: z=0/1."""
    ny = nx if ny is None else ny
    nz = nx if nz is None else nz
    xs = np.linspace(0.0, 1.0, nx + 1)
    ys = np.linspace(0.0, 1.0, ny + 1)
    zs = np.linspace(0.0, 1.0, nz + 1)
    Z, Y, X = np.meshgrid(zs, ys, xs, indexing="ij")
    coords = np.stack([X.ravel(), Y.ravel(), Z.ravel()], axis=1)
    sx, sy, sz = 1, nx + 1, (nx + 1) * (ny + 1)
    kk, jj, ii = np.meshgrid(np.arange(nz), np.arange(ny), np.arange(nx), indexing="ij")
    v0 = (kk * sz + jj * sy + ii).ravel()
    hexv = np.stack(
        [v0, v0 + sx, v0 + sy, v0 + sx + sy, v0 + sz, v0 + sx + sz, v0 + sy + sz, v0 + sx + sy + sz],
        axis=1,
    )
    cells = hexv[:, _KUHN].reshape(-1, 4).astype(np.int32)
    # boundary faces: every cell face lying in a bounding plane
    faces = np.concatenate([cells[:, [1, 2, 3]], cells[:, [0, 2, 3]], cells[:, [0, 1, 3]], cells[:, [0, 1, 2]]])
    fc = coords[faces]  # [4C,3,3]
    bf, bt = [], []
    for tag, (ax, val) in enumerate([(0, 0.0), (0, 1.0), (1, 0.0), (1, 1.0), (2, 0.0), (2, 1.0)], start=1):
        sel = np.all(fc[:, :, ax] == val, axis=1)
        bf.append(faces[sel]); bt.append(np.full(sel.sum(), tag))
    return OMesh(coords, cells, np.concatenate(bf), np.concatenate(bt))


def unit_interval_mesh(n):
    coords = np.linspace(0.0, 1.0, n + 1)[:, None]
    cells = np.stack([np.arange(n), np.arange(n) + 1], axis=1)
    return OMesh(coords, cells, np.array([[0], [n]]), np.array([1, 2]))


def _periodic_from(mesh, shape):
    """Identify opposite faces of a structured unit mesh whose vertices are numbered on the lattice `shape`
    (vertex counts per direction, first index slowest): periodic P1 map + DG coordinates."""
    idx = np.arange(int(np.prod(shape))).reshape(shape)
    per = np.arange(int(np.prod([s - 1 for s in shape]))).reshape([s - 1 for s in shape])
    wrap = per
    for ax, s in enumerate(shape):
        wrap = np.concatenate([wrap, np.take(wrap, [0], axis=ax)], axis=ax)
    g2p = np.empty(idx.size, dtype=np.int64)
    g2p[idx.ravel()] = wrap.ravel()
    cells = g2p[mesh.cells].astype(np.int32)
    coords = mesh.coords[mesh.cells].reshape(-1, mesh.dim)
    return OMesh(coords, cells, np.zeros((0, mesh.dim), np.int32), np.zeros(0, np.int32), periodic=True)


def periodic_unit_interval_mesh(n):
    return _periodic_from(unit_interval_mesh(n), (n + 1,))


def periodic_unit_square_mesh(nx, ny=None):
    """PeriodicUnitSquareMesh(nx, ny), both directions (demos/monge_ampere_periodic.py:16)."""
    ny = nx if ny is None else ny
    return _periodic_from(unit_square_mesh(nx, ny), (nx + 1, ny + 1))


def periodic_unit_cube_mesh(nx, ny=None, nz=None):
    ny = nx if ny is None else ny
    nz = nx if nz is None else nz
    return _periodic_from(unit_cube_mesh(nx, ny, nz), (nz + 1, ny + 1, nx + 1))


# --------------------------------------------------------------------------------------
# P1 element data on affine simplices
# --------------------------------------------------------------------------------------


def cell_jacobians(coords, cells):
    """J = [x1-x0 ... xd-x0] as columns -> [C,d,d]; detJ [C]."""
    x = coords[cells]  # [C,d+1,d]
    J = np.transpose(x[:, 1:, :] - x[:, :1, :], (0, 2, 1))
    return J, np.linalg.det(J)


def cell_geometry(coords, cells):
    """vol [C] and grad psi [C,d+1,d] (rows k>=1 = rows of J^-1; row 0 = -sum)."""
    d = coords.shape[1]
    J, det = cell_jacobians(coords, cells)
    Jinv = np.linalg.inv(J)
    grads = np.empty((cells.shape[0], d + 1, d))
    grads[:, 1:, :] = Jinv
    grads[:, 0, :] = -Jinv.sum(axis=1)
    return np.abs(det) / math.factorial(d), grads


def signed_detJ(coords, cells):
    """Tangling kernel arithmetic (movement/tangling.py:26-29): det of the edge matrix
    in cell-node-map vertex order, cofactor expansion along the first row, no FMA
    contraction (numpy evaluates each product/sum separately)."""
    x = coords[cells]
    e = x[:, 1:, :] - x[:, :1, :]  # rows = edge vectors
    d = coords.shape[1]
    if d == 1:
        return e[:, 0, 0].copy()
    if d == 2:
        return e[:, 0, 0] * e[:, 1, 1] - e[:, 0, 1] * e[:, 1, 0]
    a, b, c = e[:, 0, :], e[:, 1, :], e[:, 2, :]
    m0 = b[:, 1] * c[:, 2] - b[:, 2] * c[:, 1]
    m1 = b[:, 0] * c[:, 2] - b[:, 2] * c[:, 0]
    m2 = b[:, 0] * c[:, 1] - b[:, 1] * c[:, 0]
    return (a[:, 0] * m0 - a[:, 1] * m1) + a[:, 2] * m2


def assemble_mass(V, cells, vol):
    d1 = cells.shape[1]
    d = d1 - 1
    loc = (np.ones((d1, d1)) + np.eye(d1)) / ((d + 1) * (d + 2))
    rows = np.repeat(cells, d1, axis=1).ravel()
    cols = np.tile(cells, (1, d1)).ravel()
    vals = (vol[:, None, None] * loc[None]).ravel()
    return sp.csr_matrix((vals, (rows, cols)), shape=(V, V))


def assemble_stiffness(V, cells, vol, grads):
    d1 = cells.shape[1]
    loc = vol[:, None, None] * np.einsum("cia,cja->cij", grads, grads)
    rows = np.repeat(cells, d1, axis=1).ravel()
    cols = np.tile(cells, (1, d1)).ravel()
    return sp.csr_matrix((loc.ravel(), (rows, cols)), shape=(V, V))


def lumped_mass(V, cells, vol):
    d1 = cells.shape[1]
    ml = np.zeros(V)
    np.add.at(ml, cells.ravel(), np.repeat(vol / d1, d1))
    return ml


def simplex_quadrature(d, degree):
    """Collapsed-coordinate Gauss-Jacobi rule exact to `degree` on the unit simplex.
    Returns barycentric points [Q,d+1] and weights summing to 1."""
    n = degree // 2 + 1
    if d == 1:
        x, w = roots_jacobi(n, 0, 0)
        t = (x + 1) / 2
        return np.stack([1 - t, t], axis=1), w / 2
    if d == 2:
        x0, w0 = roots_jacobi(n, 1, 0)
        x1, w1 = roots_jacobi(n, 0, 0)
        s = (1 - x0) / 2  # Jacobi(1,0) weight (1-x) -> density s on [0,1]
        t = (x1 + 1) / 2
        S, T = np.meshgrid(s, t, indexing="ij")
        lam = np.stack([1 - S, S * (1 - T), S * T], axis=-1).reshape(-1, 3)
        W = np.outer(w0 / 4, w1 / 2).ravel() * 2.0  # int_0^1 s ds = 1/2 -> normalise to 1
        return lam, W
    if d == 3:
        x0, w0 = roots_jacobi(n, 2, 0)
        x1, w1 = roots_jacobi(n, 1, 0)
        x2, w2 = roots_jacobi(n, 0, 0)
        s = (1 - x0) / 2  # density s^2
        t = (1 - x1) / 2  # density t
        u = (x2 + 1) / 2
        S, T, U = np.meshgrid(s, t, u, indexing="ij")
        lam = np.stack([1 - S, S * (1 - T), S * T * (1 - U), S * T * U], axis=-1).reshape(-1, 4)
        W = (w0[:, None, None] / 8) * (w1[None, :, None] / 4) * (w2[None, None, :] / 2)
        return lam, W.ravel() * 6.0  # int s^2 ds * int t dt = 1/3 * 1/2 -> normalise to 1
    raise ValueError(d)


# --------------------------------------------------------------------------------------
# Boundary data
# --------------------------------------------------------------------------------------


def boundary_facet_geometry(mesh: OMesh):
    """For each boundary facet: adjacent cell, outward unit normal, measure."""
    d = mesh.dim
    cells, coords = mesh.cells, mesh.coords
    # map sorted facet vertex tuple -> cell
    key = {}
    d1 = d + 1
    for loc in range(d1):
        fv = np.delete(cells, loc, axis=1)
        fv = np.sort(fv, axis=1)
        for c, tup in enumerate(map(tuple, fv)):
            key.setdefault(tup, []).append((c, loc))
    F = mesh.bfacets.shape[0]
    fcell = np.empty(F, dtype=np.int64)
    normal = np.empty((F, d))
    area = np.empty(F)
    for f in range(F):
        tup = tuple(sorted(mesh.bfacets[f]))
        (c, loc), = key[tup]
        fcell[f] = c
        opp = coords[cells[c, loc]]
        fx = coords[mesh.bfacets[f]]
        if d == 1:
            n = np.sign(fx[0] - opp); a = 1.0
        elif d == 2:
            e = fx[1] - fx[0]
            n = np.array([e[1], -e[0]]); a = np.linalg.norm(e); n = n / a
        else:
            nn = np.cross(fx[1] - fx[0], fx[2] - fx[0]); a = 0.5 * np.linalg.norm(nn); n = nn / (2 * a)
        if np.dot(n, fx[0] - opp) < 0:
            n = -n
        normal[f] = n; area[f] = a
    return fcell, normal, area


# --------------------------------------------------------------------------------------
# Monitors (movement/monitor.py:129-143, demos/*.py, test/monitors.py)
# --------------------------------------------------------------------------------------


# --------------------------------------------------------------------------------------
# Solution-based monitors (movement/monitor.py:174-365)
# --------------------------------------------------------------------------------------
# The recovery operators come from `animate.recovery`, a third-party dependency that is neither vendored nor pinned by the
# reference (pyproject: "animate" from git, no version) and is absent from /root/reference: PARITY UNPINNED for this
# block.  Restated from its documented algorithm:
#   recover_gradient_l2(f, target_space)  = firedrake.project(grad(f), P1_vec): consistent-mass L2 projection
#   clement_interpolant(P0 field)         = at every vertex, the volume-weighted mean of the field over the cells that
#                                           contain the vertex (no boundary treatment: boundary=False is the default)
#   recover_hessian_clement(f), f in P1   = g = clement(grad f),  H = clement(grad g);  returns (g, H)
# The known answers of test/test_monitor.py:64-152 that need no P2 space are reproduced by tests/test_oracle_golden.py.


def recover_gradient_l2(mesh, u):
    """L2 projection of grad(u), u in P1, into P1_vec (movement/monitor.py:232-240)."""
    vol, grads = cell_geometry(mesh.coords, mesh.ccells)
    V, d = mesh.num_vertices, mesh.dim
    gK = np.einsum("cj,cja->ca", u[mesh.cells], grads)
    b = np.zeros((V, d))
    for loc in range(d + 1):
        np.add.at(b, mesh.cells[:, loc], (vol / (d + 1))[:, None] * gK)
    return spla.splu(assemble_mass(V, mesh.cells, vol).tocsc()).solve(b)


def clement_gradient(mesh, u):
    """Clement interpolant of the cellwise gradient of the P1 field(s) u [V] or [V,c]: out[v, (c,) a] = volume-weighted
    mean over the cells around v of d_a u_c."""
    vol, grads = cell_geometry(mesh.coords, mesh.ccells)
    V, d = mesh.num_vertices, mesh.dim
    one = u.ndim == 1
    uu = u.reshape(V, -1)
    gK = np.einsum("cjk,cja->cka", uu[mesh.cells], grads)          # [C, c, d]
    num = np.zeros((V,) + gK.shape[1:])
    den = np.zeros(V)
    for loc in range(d + 1):
        np.add.at(num, mesh.cells[:, loc], vol[:, None, None] * gK)
        np.add.at(den, mesh.cells[:, loc], vol)
    out = num / den[:, None, None]
    return out[:, 0, :] if one else out


def recover_hessian_clement(mesh, u):
    """(g, H) by two Clement interpolations (movement/monitor.py:286-293); H[v, i, j] = d_j g_i."""
    g = clement_gradient(mesh, u)
    return g, clement_gradient(mesh, g)


_PROJ_RULE = {1: (0.7886751345948129, 0.2113248654051871), 2: (2.0 / 3.0, 1.0 / 6.0), 3: (0.5854101966249685, 0.1381966011250105)}


def p1_evaluate(mesh, u, points, chunk=256):
    """u_h(p) for every row of `points`: brute force over all cells -- the cell whose smallest barycentric coordinate is
    largest contains the point (P1 fields are continuous, so ties do not matter).  Returns (values, smallest coordinate)."""
    x = mesh.coords[mesh.cells]                                   # [C, d+1, d]
    _, grads = cell_geometry(mesh.coords, mesh.cells)
    vals = np.empty(points.shape[0])
    lmin = np.empty(points.shape[0])
    for s0 in range(0, points.shape[0], chunk):
        p = points[s0:s0 + chunk]
        lam = np.einsum("cja,pca->pcj", grads, p[:, None, :] - x[None, :, 0, :])     # psi_j(p) - delta_j0
        lam[:, :, 0] += 1.0
        worst = lam.min(axis=2)
        best = worst.argmax(axis=1)
        rows = np.arange(p.shape[0])
        vals[s0:s0 + chunk] = np.einsum("pj,pj->p", lam[rows, best], u[mesh.cells[best]])
        lmin[s0:s0 + chunk] = worst[rows, best]
    return vals, lmin


def project_p1(src, u, tgt):
    """movement/monitor.py:191-201, `Function(fs).project(self.solution)` from the solution's mesh `src` onto `tgt` (same
    domain): Firedrake's supermesh projection is third-party and unpinned here; RESTATED as the Galerkin L2 projection
    M_t u_t = b, b_i = int psi_i^t u_s, the integral by the (d+1)-point degree-2 rule on every target cell (one
    barycentric coordinate A, the others B, equal weights).  Identity for coinciding meshes, exact for linear u."""
    d = tgt.dim
    A, B = _PROJ_RULE[d]
    W = np.full((d + 1, d + 1), B) + (A - B) * np.eye(d + 1)     # W[q, j] = lambda_j(x_q)
    xq = np.einsum("qj,cja->cqa", W, tgt.coords[tgt.cells])
    uq = p1_evaluate(src, u, xq.reshape(-1, d))[0].reshape(-1, d + 1)
    vol, _ = cell_geometry(tgt.coords, tgt.cells)
    b = np.zeros(tgt.num_vertices)
    contrib = (vol / (d + 1))[:, None] * (uq @ W)                # [C, j] = vol/(d+1) sum_q W[q,j] uq[q]
    for loc in range(d + 1):
        np.add.at(b, tgt.cells[:, loc], contrib[:, loc])
    return spla.splu(assemble_mass(tgt.num_vertices, tgt.cells, vol).tocsc()).solve(b)


def make_solution_monitor(src, u, gradient_scale_factor=None, hessian_scale_factor=None):
    """The callable a mover evaluates every iteration (movement/monge_ampere.py:505-506 with a SolutionBasedMonitorBuilder):
    x -> monitor on the mesh moved to x, the solution projected onto it first."""

    def m(x):
        tgt = OMesh(x, src.cells, src.bfacets, src.btags)
        ut = u if np.array_equal(x, src.coords) else project_p1(src, u, tgt)
        return solution_monitor(tgt, ut, gradient_scale_factor, hessian_scale_factor)

    return m


def solution_monitor(mesh, u, gradient_scale_factor=None, hessian_scale_factor=None):
    """movement/monitor.py:242-255 (gradient), :295-308 (Hessian), :346-365 (both): 1 + alpha gg / max|gg| + beta HH / max|HH|."""
    m = np.ones(mesh.num_vertices)
    if gradient_scale_factor is not None:
        g = recover_gradient_l2(mesh, u)
        gg = np.einsum("ia,ia->i", g, g)
        m = m + gradient_scale_factor * gg / np.abs(gg).max()
    if hessian_scale_factor is not None:
        H = recover_hessian_clement(mesh, u)[1]
        HH = np.einsum("iab,iab->i", H, H)
        m = m + hessian_scale_factor * HH / np.abs(HH).max()
    return m


def monitor_const(x):
    return np.ones(x.shape[0])


def make_ball_monitor(centre, radius, amplitude, width):
    """movement/monitor.py:138-143: 1 + a / cosh^2(w (|x-c|^2 - radius^2))."""
    centre = np.asarray(centre, dtype=float)

    def m(x):
        diff = x - centre
        dist = np.einsum("ia,ia->i", diff, diff)
        return 1.0 + amplitude / np.cosh(width * (dist - radius ** 2)) ** 2

    return m


def monitor_test_ring(x):
    """test/monitors.py:9-16 (note: gamma enters *unsquared*)."""
    diff = x - 0.5
    r = np.einsum("ia,ia->i", diff, diff)
    return 1.0 + 10.0 / np.cosh(200.0 * (r - 0.15)) ** 2


def monitor_helmholtz_exact(x, alpha=5.0):
    """demos/monge_ampere_helmholtz.py:54-61,115-122: 1 + alpha |HessU|_F / max."""
    w2 = 0.1 ** 2
    dx = x[:, 0] - 0.51
    dy = x[:, 1] - 0.65
    u = np.exp(-(dx * dx + dy * dy) / w2)
    uxx = u * (4 * dx * dx / w2 ** 2 - 2 / w2)
    uyy = u * (4 * dy * dy / w2 ** 2 - 2 / w2)
    uxy = u * (4 * dx * dy / w2 ** 2)
    hn = np.sqrt(uxx * uxx + uyy * uyy + 2 * uxy * uxy)
    return 1.0 + alpha * hn / hn.max()


def sinatan3_hess_sq(x):
    """H:H for u = 0.1 sin(50xz) + atan2(0.1, sin(5y) - 2xz) (demos/monge_ampere_3d.py:13-15,58-63)."""
    X, Y, Z = x[:, 0], x[:, 1], x[:, 2]
    s = 50 * X * Z
    sn, cs = np.sin(s), np.cos(s)
    H = np.zeros((x.shape[0], 3, 3))
    H[:, 0, 0] = -250 * Z * Z * sn
    H[:, 2, 2] = -250 * X * X * sn
    H[:, 0, 2] = H[:, 2, 0] = 5 * cs - 250 * X * Z * sn
    q = np.sin(5 * Y) - 2 * X * Z
    D = 0.01 + q * q
    g = -0.1 / D
    gp = 0.2 * q / (D * D)
    gq = np.stack([-2 * Z, 5 * np.cos(5 * Y), -2 * X], axis=1)
    H += gp[:, None, None] * gq[:, :, None] * gq[:, None, :]
    H[:, 1, 1] += g * (-25 * np.sin(5 * Y))
    H[:, 0, 2] += g * (-2.0)
    H[:, 2, 0] += g * (-2.0)
    return np.einsum("iab,iab->i", H, H)


def monitor_sinatan3(x, alpha=10.0):
    """demos/monge_ampere_3d.py:52-69: 1 + alpha (H:H)/max(H:H) -- no sqrt."""
    hn = sinatan3_hess_sq(x)
    return 1.0 + alpha * hn / hn.max()


# --------------------------------------------------------------------------------------
# Diagnostics (movement/mover.py:199-217)
# --------------------------------------------------------------------------------------


def volume_ratio(v):
    return v.max() / v.min()


def coefficient_of_variation(v):
    mean = v.sum() / v.size
    return math.sqrt(((v - mean) ** 2).sum() / v.size) / mean


def plural(n):
    return "s" if n != 1 else ""


class ConvergenceError(Exception):
    pass


# --------------------------------------------------------------------------------------
# Tangling checker (movement/tangling.py:10-53)
# --------------------------------------------------------------------------------------


class OracleTanglingChecker:
    def __init__(self, coords, cells, raise_error=True):
        self.cells = cells
        self.detJ0 = signed_detJ(coords, cells)
        self.raise_error = raise_error

    def ratio(self, coords):
        return signed_detJ(coords, self.cells) / self.detJ0

    def check(self, coords):
        r = self.ratio(coords)
        n = int((r <= 0).sum())
        if n > 0:
            msg = f"Mesh has {n} tangled element{'s' if n > 1 else ''}."
            if self.raise_error:
                raise ValueError(msg)
        return n


# --------------------------------------------------------------------------------------
# Monge-Ampere operators on the fixed computational mesh
# --------------------------------------------------------------------------------------


class MAOperators:
    """Everything constant on the xi-mesh: M, K, m_l, vol, grads, BC
    masks, boundary facet data, quadrature tables."""

    def __init__(self, mesh: OMesh, fixed_boundary_segments=(), solver="lu"):
        # solver: "lu" = sparse LU (exact, the default everywhere); "cg" = Jacobi-preconditioned conjugate gradients
        # driven to a 2-norm relative residual of 1e-14 -- used only by tests/golden/make_midsize_golden.py for the 3D
        # meshes whose LU fill does not fit a test run, and checked there against "lu" on a smaller mesh.
        self.solver = solver
        self.mesh = mesh
        d = self.dim = mesh.dim
        V = self.V = mesh.num_vertices
        self.xi = mesh.coords.copy()
        self.cells = mesh.cells
        self.ccells = mesh.ccells
        self.vol, self.grads = cell_geometry(self.xi, self.ccells)
        self.total_volume = self.vol.sum()
        self.M = assemble_mass(V, self.cells, self.vol)
        self.K = assemble_stiffness(V, self.cells, self.vol, self.grads)
        self.ml = lumped_mass(V, self.cells, self.vol)
        self.fcell, self.fnormal, self.farea = boundary_facet_geometry(mesh)
        self.qlam, self.qw = simplex_quadrature(d, d + 2)
        # -- L2 projector BCs, axis-aligned branch (movement/monge_ampere.py:246-272)
        tags = list(np.unique(mesh.btags))
        if fixed_boundary_segments == "on_boundary" or list(fixed_boundary_segments) == ["on_boundary"]:
            fixed_boundary_segments = tags
        for t in fixed_boundary_segments:
            if t not in tags:
                raise ValueError(f"Provided boundary_tag '{t}' is invalid.")
        mask = np.zeros((V, d), dtype=bool)  # True = constrained dof
        self.eq_tags = []                     # non-axis-aligned free segments: EquationBC branch (monge_ampere.py:274-309)
        self.bc_kinds = []                    # per tag, in tag order: "dirichlet" | "component" | "equation"
        for t in tags:
            sel = mesh.btags == t
            nodes = np.unique(mesh.bfacets[sel])
            if t in fixed_boundary_segments or d == 1:
                mask[nodes, :] = True
                self.bc_kinds.append("dirichlet")
                continue
            with np.errstate(invalid="ignore"):
                absn = (np.abs(self.fnormal[sel]) * self.farea[sel, None]).sum(axis=0)
            iszero = np.isclose(absn, 0.0)
            assert iszero.sum() < d
            if iszero.sum() == d - 1:
                mask[nodes, int(np.where(~iszero)[0][0])] = True
                self.bc_kinds.append("component")
                continue
            self.eq_tags.append(t)
            self.bc_kinds.append("equation")
            mask[nodes, :] = True             # every component of these nodes is prescribed by the boundary solve
        self.bc_mask = mask
        self.bc_error = None
        self._eq = [self._equation_bc(t) for t in self.eq_tags]
        self._Mlu = None
        self._Mbc_lu = {}
        self._Klu = None

    # -- exact solves ---------------------------------------------------------------
    @staticmethod
    def _cg(A, b, project=False):
        """Jacobi-PCG with residual replacement, column by column (project: zero-mean iterates for the singular K).
        A sweep runs until the recurrence residual is below 1e-14 ||b||_2, then the TRUE residual b - A x is formed and
        the iteration restarts from it; it ends when the true residual is below 1e-14 ||b||_2 or has stopped
        decreasing (the eps * ||A|| ||x|| floor that bounds a backward-stable direct solve as well)."""
        b = np.asarray(b, dtype=float)
        B = b.reshape(b.shape[0], -1)
        dinv = 1.0 / A.diagonal()
        X = np.zeros_like(B)
        for c in range(B.shape[1]):
            bb = B[:, c]
            bn = np.linalg.norm(bb)
            if bn == 0.0:
                continue
            x = np.zeros_like(bb)
            r = bb.copy()
            last_true = np.inf
            for _sweep in range(12):
                z = dinv * r
                if project:
                    z -= z.mean()
                p = z.copy()
                rz = r @ z
                for _ in range(10 * A.shape[0]):
                    q = A @ p
                    alpha = rz / (p @ q)
                    x += alpha * p
                    r -= alpha * q
                    if np.linalg.norm(r) <= 1e-14 * bn:
                        break
                    z = dinv * r
                    if project:
                        z -= z.mean()
                    rzn = r @ z
                    p = z + (rzn / rz) * p
                    rz = rzn
                else:
                    raise RuntimeError("oracle cg: not converged")
                r = bb - A @ x
                tn = np.linalg.norm(r)
                if tn <= 1e-14 * bn or tn >= 0.5 * last_true:
                    break
                last_true = tn
            if tn > 1e-12 * bn:
                raise RuntimeError(f"oracle cg: true residual stagnated at {tn / bn:.2e}")
            X[:, c] = x
        return X.reshape(b.shape)

    # -- EquationBC branch of the L2 projector (movement/monge_ampere.py:274-309) -----------------------------------
    def _equation_bc(self, tag):
        """Structures of one straight / planar, non-axis-aligned, free boundary segment.

        The reference imposes two EquationBCs on the segment's rows: int (v.n)(u.n) ds = 0 (bc1) and
        int v_t . u_t ds = int v_t . grad(phi_old)_t ds with the nodes shared with any other segment pinned to zero
        (bc2, its nested DirichletBC `ffacet_bc`).  Both are assembled into the same rows (Firedrake adds the forms
        of successive EquationBCs), and n n^T + (I - n n^T) = I, so the rows read int v . u ds = int v . P_t grad(phi_old) ds:
        the boundary-mass (L2(Gamma)) projection of the tangential gradient.  Its unique solution has u = 0 at the
        shared ('corner') nodes and u.n = 0 everywhere on the segment; the interior rows then see these values as
        inhomogeneous Dirichlet data.  Also performs the reference's straight-line / plane check (:288-299), whose
        error is raised at the first projection like the reference's lazily built solver does."""
        mesh, d = self.mesh, self.dim
        sel = np.where(mesh.btags == tag)[0]
        nodes = np.unique(mesh.bfacets[sel])
        others = np.unique(mesh.bfacets[mesh.btags != tag])
        corners = np.intersect1d(nodes, others)
        # hyperplane through the corner nodes (math.py:7-42; a deterministic choice of the d points)
        name = {2: "line", 3: "plane"}[d]
        pts = mesh.coords[corners]
        assert len(pts) >= d
        p0 = pts[0]
        dist0 = np.linalg.norm(pts - p0, axis=1)
        p1 = pts[int(np.argmax(dist0))]
        err = None
        if not dist0.max() > 0:
            err = ValueError(f"Could not determine a {name} for the provided points.")
        elif d == 2:
            e = (p1 - p0) / np.linalg.norm(p1 - p0)
            x = mesh.coords[nodes] - p0
            dist = np.abs(x[:, 0] * e[1] - x[:, 1] * e[0])
        else:
            cr = np.cross(p1 - p0, pts - p0)
            k = int(np.argmax(np.linalg.norm(cr, axis=1)))
            if not np.linalg.norm(cr[k]) > 1e-14 * dist0.max() ** 2:
                err = ValueError(f"Could not determine a {name} for the provided points.")
            else:
                nrm = cr[k] / np.linalg.norm(cr[k])
                dist = np.abs((mesh.coords[nodes] - p0) @ nrm)
        if err is None and not np.all(np.isclose(dist, 0)):
            err = ValueError(f"Boundary segment '{tag}' is not {'linear' if d == 2 else 'planar'}.")
        if err is not None and self.bc_error is None:
            self.bc_error = err
        unknown = np.setdiff1d(nodes, corners)
        loc = -np.ones(self.V, dtype=np.int64)
        loc[unknown] = np.arange(len(unknown))
        # boundary mass matrix of the segment over the unknown nodes: facet matrix area / (d (d+1)) (1 + delta_ab)
        fv = mesh.bfacets[sel]
        a = self.farea[sel] / (d * (d + 1))
        rows, cols, vals = [], [], []
        for i in range(d):
            for j in range(d):
                ok = (loc[fv[:, i]] >= 0) & (loc[fv[:, j]] >= 0)
                rows.append(loc[fv[ok, i]]); cols.append(loc[fv[ok, j]]); vals.append(a[ok] * (2.0 if i == j else 1.0))
        nu = len(unknown)
        Mb = sp.csr_matrix((np.concatenate(vals), (np.concatenate(rows), np.concatenate(cols))), shape=(nu, nu))
        return dict(tag=tag, facets=sel, nodes=nodes, corners=corners, unknown=unknown, loc=loc,
                    lu=spla.splu(Mb.tocsc()) if nu else None)

    def equation_bc_values(self, phi):
        """[V,d] values of grad(phi) at the nodes of the EquationBC segments (zero elsewhere)."""
        d = self.dim
        g = np.zeros((self.V, d))
        if not self._eq:
            return g
        gK = np.einsum("cj,cja->ca", phi[self.cells], self.grads)
        for eq in self._eq:
            sel = eq["facets"]
            n = self.fnormal[sel]
            gf = gK[self.fcell[sel]]
            gt = gf - n * np.einsum("fa,fa->f", n, gf)[:, None]            # tangential(grad phi, n), per facet
            contrib = (self.farea[sel] / d)[:, None] * gt                  # int_f psi_i ds = area / d
            rhs = np.zeros((len(eq["unknown"]), d))
            fv = self.mesh.bfacets[sel]
            for i in range(d):
                li = eq["loc"][fv[:, i]]
                ok = li >= 0
                np.add.at(rhs, li[ok], contrib[ok])
            if len(eq["unknown"]):
                g[eq["unknown"]] = eq["lu"].solve(rhs)
        return g

    def solve_mass(self, rhs):
        if self.solver == "cg":
            return self._cg(self.M, rhs)
        if self._Mlu is None:
            self._Mlu = spla.splu(self.M.tocsc())
        return self._Mlu.solve(rhs)

    def solve_mass_bc(self, rhs, comp):
        """M g = b with symmetric elimination of homogeneous Dirichlet dofs."""
        if comp not in self._Mbc_lu:
            free = ~self.bc_mask[:, comp]
            P = sp.diags(free.astype(float))
            A = (P @ self.M @ P + sp.diags((~free).astype(float))).tocsr()
            self._Mbc_lu[comp] = (A if self.solver == "cg" else spla.splu(A.tocsc()), free)
        lu, free = self._Mbc_lu[comp]
        if self.solver == "cg":
            return self._cg(lu, np.where(free, rhs, 0.0))
        return lu.solve(np.where(free, rhs, 0.0))

    def solve_mass_chebyshev(self, rhs, x0=None, fixed=None, rtol=1e-12, atol=1e-50, maxit=500, kmin=6, kdense=18):
        """CPU restatement of the GPU path's iterative mass solver (movement_b200/csrc/mv_cheb.cu), used to pin its
        iteration counts: Chebyshev iteration on diag(M)^-1 M over Wathen's interval [1/2, 1 + d/2] for P1
        simplices, x_{k+1} = x_k + a_k (x_k - x_{k-1}) + c_k D^-1 (b - M x_k), stopped when every component has
        sqrt(r.D^-1 r) <= max(rtol * ref, atol), ref = max(b.D^-1 b, r_0.D^-1 r_0), checked at k = 0 (a converged
        guess is returned untouched), k = kmin, kmin + 2, ... and every k >= kdense.  Replaces the KSP `cg` of
        movement/solver_parameters.py:62-66 at movement/monge_ampere.py:340,483 -- same linear systems, so
        results agree with `solve_mass` to the tolerance.  `fixed`: boolean [V, c] mask of Dirichlet dofs whose
        value in x0 is kept.  Returns (x, iterations)."""
        b = np.asarray(rhs, dtype=float)
        one = b.ndim == 1
        b = b.reshape(self.V, -1)
        x = np.zeros_like(b) if x0 is None else np.array(x0, dtype=float).reshape(self.V, -1)
        fix = np.zeros(b.shape, bool) if fixed is None else np.asarray(fixed, bool).reshape(self.V, -1)
        dinv = 1.0 / self.M.diagonal()
        lmin, lmax = 0.5, 1.0 + 0.5 * self.dim
        theta, delta = 0.5 * (lmax + lmin), 0.5 * (lmax - lmin)
        sigma1 = theta / delta
        rho = 1.0 / sigma1
        xprev = x.copy()
        thr = None
        for k in range(maxit):
            r = np.where(fix, 0.0, b - self.M @ x)
            z = dinv[:, None] * r
            check = k == 0 or k >= kdense or (k >= kmin and (k - kmin) % 2 == 0)
            if check:
                rz = (r * z).sum(axis=0)
                if k == 0:
                    bb = np.where(fix, 0.0, b * dinv[:, None] * b).sum(axis=0)
                    thr = np.maximum(rtol ** 2 * np.maximum(rz, bb), atol ** 2)
                if np.all(rz <= thr):
                    return (x[:, 0] if one else x), k
            if k == 0:
                ak, ck = 0.0, 1.0 / theta
            else:
                rk = 1.0 / (2.0 * sigma1 - rho)
                ak, ck = rk * rho, 2.0 * rk / delta
                rho = rk
            xnew = np.where(fix, x, x + ak * (x - xprev) + ck * z)
            xprev, x = x, xnew
        raise RuntimeError(f"chebyshev: not converged in {maxit} iterations")

    def solve_mass_chebyshev_mixed(self, rhs, x0=None, fixed=None, rtol=1e-12, atol=1e-50, maxit=500, sweep=16):
        """CPU restatement of the GPU path's default mass solver (movement_b200/csrc/mv_cheb.cu, cheb_mixed): the same
        Chebyshev polynomial as `solve_mass_chebyshev` used as the inner solver of an iterative refinement.  A
        residual pass in double (r = b - M x, the stopping test of `solve_mass_chebyshev` on r.D^-1 r) is followed by
        a sweep of k Chebyshev iterations on the correction M d = r, d_0 = 0, carried out in single precision
        (float32 copies of M, D^-1, r and the iterates), and x += d_k in double; k = clamp(ceil(log(need / 4) /
        log(fac)), 2, sweep) with need = min_c sqrt(thr_c / rz_c) and fac = (sqrt(kappa) - 1) / (sqrt(kappa) + 1).
        Same linear systems as movement/monge_ampere.py:340,483 (KSP `cg`, movement/solver_parameters.py:62-66), so
        results agree with `solve_mass` to the tolerance.  Returns (x, iterations = single-precision iterations +
        residual passes after the first, residual passes)."""
        b = np.asarray(rhs, dtype=float)
        one = b.ndim == 1
        b = b.reshape(self.V, -1)
        x = np.zeros_like(b) if x0 is None else np.array(x0, dtype=float).reshape(self.V, -1)
        fix = np.zeros(b.shape, bool) if fixed is None else np.asarray(fixed, bool).reshape(self.V, -1)
        dinv = 1.0 / self.M.diagonal()
        M32 = self.M.astype(np.float32)
        dinv32 = dinv.astype(np.float32)[:, None]
        lmin, lmax = 0.5, 1.0 + 0.5 * self.dim
        theta, delta = 0.5 * (lmax + lmin), 0.5 * (lmax - lmin)
        sigma1 = theta / delta
        fac = (np.sqrt(lmax / lmin) - 1.0) / (np.sqrt(lmax / lmin) + 1.0)
        its, passes, thr = 0, 0, None
        while True:
            r = np.where(fix, 0.0, b - self.M @ x)
            z = dinv[:, None] * r
            rz = (r * z).sum(axis=0)
            passes += 1
            if thr is None:
                bb = np.where(fix, 0.0, b * dinv[:, None] * b).sum(axis=0)
                thr = np.maximum(rtol ** 2 * np.maximum(rz, bb), atol ** 2)
            if np.all(rz <= thr):
                return (x[:, 0] if one else x), its + passes - 1, passes
            if its + passes - 1 >= maxit:
                raise RuntimeError(f"chebyshev: not converged in {maxit} iterations")
            live = (rz > thr) & (rz > 0)
            need = min(1.0, float(np.sqrt(thr[live] / rz[live]).min()))
            k = int(min(max(np.ceil(np.log(0.25 * need) / np.log(fac)), 2), max(sweep, 2)))
            r32 = r.astype(np.float32)
            d = (z / theta).astype(np.float32)           # d_1, computed in double by the residual pass
            dprev = np.zeros_like(d)
            rho = 1.0 / sigma1
            for _ in range(1, k):
                rk = 1.0 / (2.0 * sigma1 - rho)
                ak, ck = np.float32(rk * rho), np.float32(2.0 * rk / delta)
                rho = rk
                res = np.where(fix, np.float32(0), r32 - M32 @ d)
                dn = np.where(fix, np.float32(0), d + ak * (d - dprev) + ck * (dinv32 * res)).astype(np.float32)
                dprev, d = d, dn
            x = x + d.astype(float)
            its += k

    def solve_poisson_nullspace(self, rhs):
        """K u = rhs with constant nullspace: rhs and u projected to zero mean."""
        V = self.V
        if self.solver == "cg":
            u = self._cg(self.K, rhs - rhs.mean(), project=True)
            return u - u.mean()
        if self._Klu is None:
            A = sp.bmat([[self.K, np.ones((V, 1))], [np.ones((1, V)), None]], format="csc")
            self._Klu = spla.splu(A)
        r = rhs - rhs.mean()
        return self._Klu.solve(np.append(r, 0.0))[:V]

    # -- A1: L2 projection of grad(phi)  (movement/monge_ampere.py:311-342) ---------
    def l2_rhs(self, phi):
        gK = np.einsum("cj,cja->ca", phi[self.cells], self.grads)  # [C,d]
        b = np.zeros((self.V, self.dim))
        contrib = (self.vol / (self.dim + 1))[:, None] * gK
        for loc in range(self.dim + 1):
            np.add.at(b, self.cells[:, loc], contrib)
        return b

    def l2_project_grad(self, phi):
        if self.bc_error is not None:
            raise self.bc_error
        b = self.l2_rhs(phi)
        if not self._eq:
            return np.stack([self.solve_mass_bc(b[:, a], a) for a in range(self.dim)], axis=1)
        # EquationBC segments: their nodes carry the boundary solve's values, the remaining rows see them as data
        gB = self.equation_bc_values(phi)
        b = b - self.M @ gB
        g = np.stack([self.solve_mass_bc(b[:, a], a) for a in range(self.dim)], axis=1)
        return g + gB

    # -- A5: theta and residual  (movement/monge_ampere.py:405-410,514,517,212-223) --
    def assemble_a(self, m, H):
        """a_i = int psi_i m det(I + H) dxi, exact (degree d+2) quadrature."""
        d = self.dim
        mq = np.einsum("qj,cj->cq", self.qlam, m[self.cells])
        Hq = np.einsum("qj,cjab->cqab", self.qlam, H[self.cells]) + np.eye(d)
        f = mq * np.linalg.det(Hq) * self.qw[None, :] * self.vol[:, None]  # [C,Q]
        loc = np.einsum("cq,qj->cj", f, self.qlam)
        a = np.zeros(self.V)
        np.add.at(a, self.cells.ravel(), loc.ravel())
        return a

    def theta_residual(self, m, H):
        a = self.assemble_a(m, H)
        theta = a.sum() / self.total_volume
        r = a - theta * self.ml
        return a, theta, r, np.linalg.norm(r) / np.linalg.norm(theta * self.ml)

    # -- A8: Hessian recovery RHS  (movement/monge_ampere.py:451-485) --------------
    def hess_rhs(self, phi):
        d = self.dim
        gK = np.einsum("cj,cja->ca", phi[self.cells], self.grads)
        c = np.zeros((self.V, d, d))
        # interior: -vol * d_b psi_i * g_a
        for loc in range(d + 1):
            np.add.at(c, self.cells[:, loc], -self.vol[:, None, None] * gK[:, :, None] * self.grads[:, loc, None, :])
        # boundary: area/d * t_a n_b,  t = g - (g.n) n
        g = gK[self.fcell]
        n = self.fnormal
        t = g - np.einsum("fa,fa->f", g, n)[:, None] * n
        fb = (self.farea / d)[:, None, None] * t[:, :, None] * n[:, None, :]
        for loc in range(d):
            np.add.at(c, self.mesh.bfacets[:, loc], fb)
        return c

    def recover_hessian(self, phi):
        c = self.hess_rhs(phi)
        d = self.dim
        return self.solve_mass(c.reshape(self.V, d * d)).reshape(self.V, d, d)

    # -- A4: volume diagnostic  (movement/monge_ampere.py:507-510) -----------------
    def monitor_volumes(self, x, m, vol0):
        volx, _ = cell_geometry(x, self.ccells)
        return volx * m[self.cells].mean(axis=1) / vol0


def galerkin_guess(xs, bs, b, tol=1e-10):
    """Initial guess of K x = b from earlier solutions xs[i] with right-hand sides bs[i] = K xs[i] (newest first): the
    energy-optimal combination sum_i a_i xs[i], G a = f with G_ij = xs[i].bs[j], f_i = xs[i].b, by a Cholesky factorisation
    in the given order that stops at the first negligible pivot (restatement of k_hist_solve / mv_ma_pseudotimestep in
    movement_b200/csrc/mv_ma.cu, which prepares the KSP solve of movement/monge_ampere.py:441-448; the reference itself
    starts every solve from zero).  Returns (x0, number of vectors used)."""
    m = len(xs)
    if m == 0:
        return np.zeros_like(b), 0
    G = np.array([[float(np.dot(xs[i], bs[j])) for j in range(m)] for i in range(m)])
    f = np.array([float(np.dot(xs[i], b)) for i in range(m)])
    L = np.zeros((m, m))
    mm = 0
    for j in range(m):
        d = G[j, j] - float(np.dot(L[j, :j], L[j, :j]))
        if not (G[j, j] > 0.0 and d > tol * G[j, j]):
            break
        L[j, j] = np.sqrt(d)
        for i in range(j + 1, m):
            L[i, j] = (G[i, j] - float(np.dot(L[i, :j], L[j, :j]))) / L[j, j]
        mm = j + 1
    if mm == 0:
        return np.zeros_like(b), 0
    y = np.linalg.solve(L[:mm, :mm], f[:mm])
    a = np.linalg.solve(L[:mm, :mm].T, y)
    return sum(a[i] * xs[i] for i in range(mm)), mm


class OracleRelaxation:
    """movement/monge_ampere.py:345-540 with exact inner solves."""

    def __init__(self, mesh: OMesh, monitor, pseudo_timestep=0.1, maxiter=1000, rtol=1e-8, dtol=2.0,
                 fixed_boundary_segments=(), raise_convergence_errors=True, verbose=False, solver="lu"):
        self.ops = MAOperators(mesh, fixed_boundary_segments, solver=solver)
        self.monitor = monitor
        self.dt = pseudo_timestep
        self.maxiter, self.rtol, self.dtol = maxiter, rtol, dtol
        self.raise_errors = raise_convergence_errors
        d, V = self.ops.dim, self.ops.V
        self.xi = self.ops.xi
        self.x = self.xi.copy()
        self.phi_old = np.zeros(V)
        self.H_old = np.zeros((V, d, d))
        self.phi = np.zeros(V)
        self.H = np.zeros((V, d, d))
        self.vol0 = self.ops.vol.copy()
        self.tangling = OracleTanglingChecker(self.xi, self.ops.ccells, raise_error=raise_convergence_errors)
        self.trace = []
        self.verbose = verbose

    def _fail(self, msg):
        raise (ConvergenceError if self.raise_errors else Warning)(msg)

    def move(self):
        o = self.ops
        i = 0
        for i in range(self.maxiter):
            self.grad_phi = o.l2_project_grad(self.phi_old)
            self.x = self.xi + o.mesh.to_coordinate_nodes(self.grad_phi)   # CG -> DG on periodic meshes (:236-239)
            self.tangling.check(self.xi)  # quirk: mesh.coordinates holds xi here
            self.m = self.monitor(o.mesh.at_vertices(self.x))
            self.volume = o.monitor_volumes(self.x, self.m, self.vol0)
            _, self.theta, r, residual = o.theta_residual(self.m, self.H_old)
            if i == 0:
                initial_norm = residual
            line = (i, volume_ratio(self.volume), coefficient_of_variation(self.volume), residual)
            self.trace.append(line)
            if self.verbose:
                print(f"{i:4d}   Volume ratio {line[1]:5.2f}   Variation (σ/μ) {line[2]:8.2e}   Residual {residual:8.2e}")
            if residual < self.rtol:
                break
            if residual > self.dtol * initial_norm:
                self._fail(f"Solver diverged after {i + 1} iteration{plural(i + 1)}.")
            if i == self.maxiter - 1:
                self._fail(f"Solver failed to converge in {i + 1} iteration{plural(i + 1)}.")
            self.phi = self.phi_old + o.solve_poisson_nullspace(self.dt * r)
            self.H = o.recover_hessian(self.phi)
            self.phi_old = self.phi.copy()
            self.H_old = self.H.copy()
        return i


class OracleQuasiNewton:
    """movement/monge_ampere.py:543-750 + PETSc newtonls / l2 line search as restated in
    SURVEY §8a-QN (exact Jacobian solve, mean-zero constraint on delta-phi)."""

    def __init__(self, mesh: OMesh, monitor, maxiter=1000, rtol=1e-8, fixed_boundary_segments=(), verbose=False,
                 snes_rtol=1e-8):
        self.ops = MAOperators(mesh, fixed_boundary_segments)
        self.monitor = monitor
        self.maxiter, self.rtol = maxiter, rtol
        self.snes_rtol = snes_rtol  # solver_parameters.py:15
        d, V = self.ops.dim, self.ops.V
        self.xi = self.ops.xi
        self.x = self.xi.copy()
        self.vol0 = self.ops.vol.copy()
        self.verbose = verbose
        self.trace = []
        self.nfunc = 0
        self._build_S()

    def _build_S(self):
        """S_ab: linear map phi -> hess_rhs(phi)_ab as sparse matrices."""
        o = self.ops
        d, V = o.dim, o.V
        cells = o.cells
        self.S = [[None] * d for _ in range(d)]
        d1 = d + 1
        n = o.fnormal
        for a in range(d):
            for b in range(d):
                rows, cols, vals = [], [], []
                for li in range(d1):
                    for lj in range(d1):
                        rows.append(cells[:, li]); cols.append(cells[:, lj])
                        vals.append(-o.vol * o.grads[:, li, b] * o.grads[:, lj, a])
                # boundary: area/d * (g_a - (g.n) n_a) n_b with g = sum_j phi_j grad_j(fcell)
                gr = o.grads[o.fcell]  # [F,d1,d]
                tcoef = gr[:, :, a] - np.einsum("fjc,fc->fj", gr, n) * n[:, None, a]  # [F,d1]
                for lf in range(d):
                    for lj in range(d1):
                        rows.append(o.mesh.bfacets[:, lf]); cols.append(cells[o.fcell, lj])
                        vals.append(o.farea / d * tcoef[:, lj] * n[:, b])
                self.S[a][b] = sp.csr_matrix(
                    (np.concatenate(vals), (np.concatenate(rows), np.concatenate(cols))), shape=(V, V)
                )

    def update_monitor(self, phi, H):
        o = self.ops
        g = o.l2_project_grad(phi)
        self.x = self.xi + o.mesh.to_coordinate_nodes(g)
        self.m = self.monitor(o.mesh.at_vertices(self.x))
        a = o.assemble_a(self.m, H)
        self.theta = a.sum() / o.total_volume
        return a

    def func(self, u):
        o = self.ops
        d, V = o.dim, o.V
        phi, H = u[:V], u[V:].reshape(V, d, d)
        a = self.update_monitor(phi, H)
        self.nfunc += 1
        Fpsi = -(a - self.theta * o.ml)
        Ftau = np.empty((V, d, d))
        for aa in range(d):
            for bb in range(d):
                Ftau[:, aa, bb] = o.M @ H[:, aa, bb] - self.S[aa][bb] @ phi
        return np.concatenate([Fpsi, Ftau.ravel()])

    def jacobian(self, u):
        """J = [[0, -A_ab],[-S_ab, M]] with m, theta frozen."""
        o = self.ops
        d, V = o.dim, o.V
        H = u[V:].reshape(V, d, d)
        cells = o.cells
        d1 = d + 1
        Hq = np.einsum("qj,cjab->cqab", o.qlam, H[cells]) + np.eye(d)
        det = np.linalg.det(Hq)
        cof = det[..., None, None] * np.transpose(np.linalg.inv(Hq), (0, 1, 3, 2))  # cofactor matrix
        mq = np.einsum("qj,cj->cq", o.qlam, self.m[cells])
        wq = mq * o.qw[None, :] * o.vol[:, None]
        blocks = [[None] * (1 + d * d) for _ in range(1 + d * d)]
        rows = np.repeat(cells, d1, axis=1).ravel()
        cols = np.tile(cells, (1, d1)).ravel()
        for a in range(d):
            for b in range(d):
                loc = np.einsum("cq,qi,qj->cij", wq * cof[:, :, a, b], o.qlam, o.qlam)
                A = sp.csr_matrix((loc.ravel(), (rows, cols)), shape=(V, V))
                k = 1 + a * d + b
                blocks[0][k] = -A
                blocks[k][0] = -self.S[a][b]
                blocks[k][k] = o.M
        # dof ordering: u = [phi (V), H flattened (V,d,d) -> index i*d*d + a*d + b]
        Jblk = sp.bmat(blocks, format="csr")
        # permutation from block ordering [phi, H_00(V), H_01(V), ...] to u ordering
        perm = np.concatenate([np.arange(V)] + [V + np.arange(V) * d * d + k for k in range(d * d)])
        P = sp.csr_matrix((np.ones(perm.size), (perm, np.arange(perm.size))), shape=(perm.size, perm.size))
        return (P @ Jblk @ P.T).tocsc()

    def solve_newton(self, J, F):
        V = self.ops.V
        N = J.shape[0]
        e = np.zeros((N, 1)); e[:V] = 1.0
        A = sp.bmat([[J, e], [e.T, None]], format="csc")
        return spla.splu(A).solve(np.append(F, 0.0))[:N]

    def l2_linesearch(self, u, Y, fnorm2):
        lam, lam_old = 0.8, 0.0
        f_old = fnorm2
        lam_mid = 0.5 * (lam + lam_old)
        for _ in range(5):
            f_mid = np.sum(self.func(u - lam_mid * Y) ** 2)
            f = np.sum(self.func(u - lam * Y) ** 2)
            delta = lam - lam_old
            d1 = (3 * f - 4 * f_mid + f_old) / delta
            d1_old = (-3 * f_old + 4 * f_mid - f) / delta
            d2 = (d1 - d1_old) / delta
            if d2 > 0:
                lam_new = lam - d1 / d2
            elif d2 < 0:
                lam_new = lam + d1 / d2
            else:
                break
            if lam_new < 1e-12:
                lam_new = 0.5 * (lam + lam_old)
            if not np.isfinite(lam_new) or lam_new > 1.05:
                break
            lam_old, lam, f_old = lam, lam_new, f
            lam_mid = 0.5 * (lam + lam_old)
        return lam

    def diagnostics(self, u):
        o = self.ops
        d, V = o.dim, o.V
        vol = o.monitor_volumes(self.x, self.m, self.vol0)
        _, _, _, res = o.theta_residual(self.m, u[V:].reshape(V, d, d))
        return volume_ratio(vol), coefficient_of_variation(vol), res

    def move(self):
        o = self.ops
        d, V = o.dim, o.V
        u = np.zeros(V * (1 + d * d))
        F = self.func(u)
        fnorm0 = np.linalg.norm(F)
        it = 0
        while True:
            fnorm = np.linalg.norm(F)
            line = (it,) + self.diagnostics(u)
            self.trace.append(line)
            if self.verbose:
                print(f"{it:4d}   Volume ratio {line[1]:5.2f}   Variation (σ/μ) {line[2]:8.2e}   Residual {line[3]:8.2e}")
            if fnorm < self.rtol or fnorm < self.snes_rtol * fnorm0:
                break
            if it == self.maxiter:
                raise ConvergenceError(f"Solver failed to converge in {it} iteration{plural(it)}.")
            J = self.jacobian(u)
            Y = self.solve_newton(J, F)
            lam = self.l2_linesearch(u, Y, fnorm ** 2)
            u = u - lam * Y
            F = self.func(u)
            it += 1
        self.phi = u[:V]
        self.H = u[V:].reshape(V, d, d)
        return it


# --------------------------------------------------------------------------------------
# Laplacian smoothing (movement/laplacian_smoothing.py:55-127)
# --------------------------------------------------------------------------------------


class OracleLaplacianSmoother:
    def __init__(self, mesh: OMesh, timestep):
        self.mesh = mesh
        self.dt = timestep
        self.xi = mesh.coords.copy()
        self.x = mesh.coords.copy()
        self.cells = mesh.cells
        vol, grads = cell_geometry(self.xi, self.cells)
        self.K = assemble_stiffness(mesh.num_vertices, self.cells, vol, grads)
        self.tangling = OracleTanglingChecker(self.xi, self.cells)
        self._lu = None

    def move(self, bc_nodes, bc_values):
        """bc_nodes: constrained vertices (all components); bc_values [len,d] velocity."""
        V, d = self.x.shape
        if self._lu is None or not np.array_equal(self._nodes, bc_nodes):
            free = np.ones(V, dtype=bool); free[bc_nodes] = False
            self._free = free
            self._nodes = np.array(bc_nodes)
            self._lu = spla.splu(self.K[free][:, free].tocsc())
            self._Kfb = self.K[free][:, ~free]
        g = np.zeros((V, d)); g[bc_nodes] = bc_values
        v = g.copy()
        v[self._free] = self._lu.solve(-self._Kfb @ g[~self._free])
        self.v = v
        self.displacement = v * self.dt
        self.x = self.x + self.displacement
        self.volume, _ = cell_geometry(self.x, self.cells)
        self.tangling.check(self.x)
        return volume_ratio(self.volume), coefficient_of_variation(self.volume), np.linalg.norm(self.displacement)


# --------------------------------------------------------------------------------------
# Lineal springs (movement/spring.py:70-283)
# --------------------------------------------------------------------------------------


def mesh_edges_2d(cells):
    """Unique edges [E,2] (i<j) and for each edge the '+' cell = lowest incident cell
    index, plus the local opposite vertex in that cell."""
    C = cells.shape[0]
    e = np.concatenate([cells[:, [1, 2]], cells[:, [0, 2]], cells[:, [0, 1]]])
    opp = np.concatenate([cells[:, 0], cells[:, 1], cells[:, 2]])
    cid = np.tile(np.arange(C), 3)
    es = np.sort(e, axis=1)
    order = np.lexsort((cid, es[:, 1], es[:, 0]))
    es, opp, cid = es[order], opp[order], cid[order]
    first = np.ones(len(es), dtype=bool)
    first[1:] = np.any(es[1:] != es[:-1], axis=1)
    return es[first], cid[first], opp[first]


def spring_edge_geometry(coords, edges, opp):
    """length, unit tangent perp(n_plus) = (-n_y, n_x), angle = arccos(clamp(t_x))
    (movement/spring.py:70-146)."""
    a, b = coords[edges[:, 0]], coords[edges[:, 1]]
    e = b - a
    L = np.linalg.norm(e, axis=1)
    n = np.stack([e[:, 1], -e[:, 0]], axis=1) / L[:, None]
    flip = np.einsum("ia,ia->i", n, a - coords[opp]) < 0
    n[flip] *= -1
    t = np.stack([-n[:, 1], n[:, 0]], axis=1)
    angle = np.arccos(np.clip(t[:, 0], -1.0, 1.0))
    return L, t, angle


def spring_stiffness(V, edges, L, angle):
    """movement/spring.py:148-181 as a sparse 2V x 2V matrix."""
    c, s = np.cos(angle), np.sin(angle)
    c2, sc, s2 = c * c / L, s * c / L, s * s / L
    i, j = edges[:, 0], edges[:, 1]
    rows, cols, vals = [], [], []
    for (p, q, sign) in [(i, i, 1), (j, j, 1), (i, j, -1), (j, i, -1)]:
        for (dr, dc, v) in [(0, 0, c2), (0, 1, sc), (1, 0, sc), (1, 1, s2)]:
            rows.append(2 * p + dr); cols.append(2 * q + dc); vals.append(sign * v)
    return sp.csr_matrix((np.concatenate(vals), (np.concatenate(rows), np.concatenate(cols))), shape=(2 * V, 2 * V))


class OracleSpringMover:
    def __init__(self, mesh: OMesh, timestep):
        self.mesh = mesh
        self.dt = timestep
        self.x = mesh.coords.copy()
        self.cells = mesh.cells
        self.edges, self.pluscell, self.opp = mesh_edges_2d(self.cells)
        self.forcing = np.zeros_like(self.x)
        self.tangling = OracleTanglingChecker(self.x, self.cells)

    def stiffness(self):
        L, t, angle = spring_edge_geometry(self.x, self.edges, self.opp)
        return spring_stiffness(self.x.shape[0], self.edges, L, angle)

    def assemble(self, bc_nodes, bc_values):
        K = self.stiffness().tolil()
        self.forcing[bc_nodes] = bc_values
        for v in bc_nodes:
            for k in (2 * v, 2 * v + 1):
                K.rows[k] = [k]; K.data[k] = [1.0]
        return K.tocsr()

    def move(self, bc_nodes, bc_values):
        K = self.assemble(bc_nodes, bc_values)
        u = spla.spsolve(K.tocsc(), self.forcing.ravel())
        self.displacement = u * self.dt
        self.x = self.x + self.displacement.reshape(self.x.shape)
        self.volume, _ = cell_geometry(self.x, self.cells)
        self.tangling.check(self.x)
        return volume_ratio(self.volume), coefficient_of_variation(self.volume), np.linalg.norm(self.displacement)
