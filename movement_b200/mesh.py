"""Mesh adaptor: the minimal Firedrake-shaped surface the movers ingest.

The reference reads ``mesh.coordinates.dat.data_with_halos`` ([V,d] f64), the P1 cell-node map
([C,d+1] i32) and the exterior-facet markers (movement/mover.py:49,83; movement/spring.py:
210-216).  On this box there is no Firedrake, so the same data comes from the generators
below, which reproduce Firedrake's ``UnitSquareMesh`` ("left" diagonal) / ``UnitCubeMesh``
(six Kuhn tetrahedra per hexahedron) layouts; :func:`from_firedrake` ingests a
live Firedrake mesh when one is importable.

A coordinate / boundary-value field is a :class:`Function`: a torch tensor (CUDA when a
device is visible, the product path; CPU otherwise so that host logic can be unit-tested)
plus a Firedrake-like ``.dat.data`` host view.
"""
from __future__ import annotations

import os

import numpy as np
import torch


def _default_device():
    if torch.cuda.is_available():
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device("cpu")


class _Dat:
    """``Function.dat``: host views of the field (``data``, ``data_ro``, ``data_with_halos``)."""

    def __init__(self, f):
        self._f = f

    @property
    def data(self):
        return self._f._host_view(writable=True)

    data_with_halos = data

    @property
    def data_ro(self):
        return self._f._host_view(writable=False)

    data_ro_with_halos = data_ro

    @property
    def norm(self):
        return float(torch.linalg.vector_norm(self._f.tensor))


class Function:
    """A nodal (P1 scalar/vector/tensor) or cellwise (P0) field.

    ``.tensor`` is the device tensor the kernels read and write.  ``.dat.data`` is a host
    numpy array: on a CUDA field it is a mirror that is refreshed from the device when the
    device copy is newer and uploaded again before the next device use (it is assumed to
    have been written once handed out, as Firedrake assumes for ``dat.data``).
    """

    def __init__(self, space_or_function, name=None, val=None):
        if isinstance(space_or_function, Function):
            src = space_or_function
            self.space = src.space
            self._t = src.tensor.clone()
        else:
            self.space = space_or_function
            self._t = torch.zeros(self.space.shape, dtype=torch.float64, device=self.space.mesh.device)
            if val is not None:
                self.assign(val)
        self.name = name
        self._host = None
        self._host_dirty = False   # host mirror may hold user writes
        self._dev_newer = True     # device copy changed since the last download
        self.dat = _Dat(self)

    # -- device side ---------------------------------------------------------------------
    @property
    def tensor(self):
        if self._host_dirty:
            self._t.copy_(self._host_t, non_blocking=False)
            self._host_dirty = False
        return self._t

    def written(self):
        """Tell the field its device tensor was modified by a kernel."""
        self._dev_newer = True
        return self

    # -- host side -----------------------------------------------------------------------
    def _host_view(self, writable):
        if self._t.device.type == "cpu":
            a = self._t.numpy()
            if not writable:
                a = a.view()
                a.flags.writeable = False
            return a
        if self._host is None:
            # page-locked mirror: refreshed and uploaded by DMA (25 GB/s) instead of through a pageable staging copy
            self._host_t = torch.empty(tuple(self._t.shape), dtype=torch.float64, pin_memory=True)
            self._host = self._host_t.numpy()
            self._dev_newer = True
        if self._dev_newer and not self._host_dirty:
            self._host_t.copy_(self._t)
            self._dev_newer = False
        if writable:
            self._host_dirty = True
            return self._host
        a = self._host.view()
        a.flags.writeable = False
        return a

    # -- Firedrake-like helpers ----------------------------------------------------------
    def function_space(self):
        return self.space

    def assign(self, other):
        t = self.tensor
        if isinstance(other, Function):
            t.copy_(other.tensor)
        elif isinstance(other, torch.Tensor):
            t.copy_(other.to(t.device, torch.float64).expand_as(t))
        else:
            a = np.asarray(other, dtype=np.float64)
            if a.shape == tuple(t.shape) and a.flags.c_contiguous and a.flags.writeable:
                t.copy_(torch.from_numpy(a))      # straight from the caller's buffer (DMA if it is pinned)
            else:
                t.copy_(torch.from_numpy(np.broadcast_to(a, tuple(t.shape)).copy()).to(t.device))
        return self.written()

    def interpolate(self, expr):
        """``expr``: Function / array / constant / callable ``x[V,d] -> values`` (vectorised
        torch).  Stands in for Firedrake's ``interpolate`` of a UFL expression."""
        if callable(expr) and not isinstance(expr, (Function, torch.Tensor)):
            val = expr(self.space.mesh.coordinates.tensor)
            if isinstance(val, (tuple, list)):
                val = torch.stack([torch.as_tensor(v, dtype=torch.float64, device=self._t.device).expand(self._t.shape[0])
                                   for v in val], dim=1)
            return self.assign(val)
        return self.assign(expr)

    def copy(self, deepcopy=True):
        return Function(self)

    def __repr__(self):
        return f"Function({self.name or ''}, shape={tuple(self._t.shape)}, device={self._t.device})"


class FunctionSpace:
    """Token naming a field layout on a mesh: 'P1', 'P1_vec', 'P1_ten', 'P0', 'edge', 'edge_vec'."""

    def __init__(self, mesh, kind):
        self.mesh = mesh
        self.kind = kind

    @property
    def shape(self):
        m, d = self.mesh, self.mesh.geometric_dimension
        return {
            "P1": (m.num_vertices(),), "P1_vec": (m.num_vertices(), d), "P1_ten": (m.num_vertices(), d, d),
            "P0": (m.num_cells(),), "DG1_vec": (m.num_cells() * (d + 1), d),
        }[self.kind]

    def __eq__(self, other):
        return isinstance(other, FunctionSpace) and other.mesh is self.mesh and other.kind == self.kind

    def __ne__(self, other):
        return not self.__eq__(other)

    def __hash__(self):
        return hash((id(self.mesh), self.kind))


class _ExteriorFacets:
    def __init__(self, mesh):
        self._mesh = mesh

    @property
    def unique_markers(self):
        return [int(t) for t in np.unique(self._mesh.btags)]


class DirichletBC:
    """Stand-in for ``firedrake.DirichletBC(space, value, tags)`` on the coordinate space
    (movement/laplacian_smoothing.py:103-106, movement/spring.py:193-221)."""

    def __init__(self, space, value, sub_domain):
        self._space = space
        self._original_arg = value
        if isinstance(sub_domain, (str, int, np.integer)):
            sub_domain = (sub_domain,)
        self.sub_domain = tuple(sub_domain)
        mesh = space.mesh
        tags = mesh.exterior_facets.unique_markers if self.sub_domain == ("on_boundary",) else self.sub_domain
        sel = np.isin(mesh.btags, np.asarray(list(tags), dtype=np.int64)) if len(tags) else np.zeros(len(mesh.btags), bool)
        self.nodes = np.unique(mesh.bfacets[sel]).astype(np.int64)

    def function_space(self):
        return self._space

    def values(self):
        """Boundary values as a [V,d] device tensor (only rows ``nodes`` are meaningful)."""
        v = self._original_arg
        shape = self._space.shape
        dev = self._space.mesh.device
        if isinstance(v, Function):
            return v.tensor
        if isinstance(v, torch.Tensor):
            return v.to(dev, torch.float64).expand(shape)
        a = np.broadcast_to(np.asarray(v, dtype=np.float64), shape)
        return torch.from_numpy(a.copy()).to(dev)


class EquationBC:
    """Descriptor of one of the two ``firedrake.EquationBC``s the reference puts on a straight, non-axis-aligned, free
    boundary segment (movement/monge_ampere.py:274-309): ``part`` is "normal" (bc1) or "tangential" (bc2).  The
    condition itself is imposed by the library (csrc/mv_eqbc.cu)."""

    def __init__(self, space, sub_domain, part):
        self._space = space
        self.sub_domain = (sub_domain,)
        self.part = part

    def function_space(self):
        return self._space


class Mesh:
    """P1 simplex mesh: device-resident coordinates + host topology.

    ``Mesh(coords, cells, bfacets, btags)`` builds from arrays; ``Mesh(function)`` (a coordinate
    :class:`Function`) makes a mesh that shares topology with the function's mesh and owns that
    function as its coordinates -- the idiom of movement/mover.py:49.
    """

    def __init__(self, coords, cells=None, bfacets=None, btags=None, device=None, part=None, degree=1, periodic=False):
        if isinstance(coords, Function):
            src = coords.space.mesh
            self.cells, self.bfacets, self.btags = src.cells, src.bfacets, src.btags
            self.device, self.part, self._degree = src.device, src.part, src._degree
            self._dim = src._dim
            self._nv = src._nv
            self.periodic = src.periodic
            self._spaces = {}
            self.coordinates = Function(self.function_space(src.coordinates.space.kind), name="Coordinates")
            self.coordinates.assign(coords)
            return
        coords = np.ascontiguousarray(coords, dtype=np.float64)
        if coords.ndim == 1:
            coords = coords[:, None]
        self._dim = coords.shape[1]
        self.cells = np.ascontiguousarray(cells, dtype=np.int32)
        # periodic: `cells` is the periodic P1 cell-node map and the coordinate field is DG1, one copy of every corner
        # per cell (row c*(d+1)+j = corner j of cell c) -- Firedrake's layout (movement/mover.py:70-75)
        self.periodic = bool(periodic)
        if self.periodic:
            assert coords.shape[0] == self.cells.shape[0] * self.cells.shape[1], "DG coordinates: one row per cell corner"
            self._nv = int(self.cells.max()) + 1
        else:
            self._nv = int(coords.shape[0])
        d = self._dim
        self.bfacets = np.zeros((0, d), np.int32) if bfacets is None else np.ascontiguousarray(bfacets, dtype=np.int32).reshape(-1, d)
        self.btags = np.zeros((0,), np.int32) if btags is None else np.ascontiguousarray(btags, dtype=np.int32)
        self.device = torch.device(device) if device is not None else _default_device()
        self.part = part          # None on one GPU; a partition.LocalPart when distributed
        self._degree = degree
        self._spaces = {}
        self.coordinates = Function(self.function_space("DG1_vec" if self.periodic else "P1_vec"), name="Coordinates")
        self.coordinates.assign(coords)

    def vertex_coordinates(self):
        """One position per P1 vertex as a device tensor [V,d]: the coordinate field itself, or, on a periodic mesh, the
        copy held by the last cell that contains the vertex (what a nodal monitor is evaluated at)."""
        if not self.periodic:
            return self.coordinates.tensor
        if not hasattr(self, "_rep"):
            rep = np.zeros(self._nv, dtype=np.int64)
            rep[self.cells.ravel()] = np.arange(self.cells.size)
            self._rep = torch.as_tensor(rep, device=self.device)
        return self.coordinates.tensor[self._rep]

    # -- Firedrake-like queries ----------------------------------------------------------
    @property
    def topological_dimension(self):
        return self._dim

    @property
    def geometric_dimension(self):
        return self._dim

    def num_vertices(self):
        return self._nv

    def num_cells(self):
        return int(self.cells.shape[0])

    @property
    def exterior_facets(self):
        return _ExteriorFacets(self)

    def function_space(self, kind):
        if kind not in self._spaces:
            self._spaces[kind] = FunctionSpace(self, kind)
        return self._spaces[kind]

    def coordinate_degree(self):
        return self._degree

    # owned counts (== totals on one GPU)
    @property
    def n_owned_vertices(self):
        return self.part.n_owned_vertices if self.part is not None else self.num_vertices()

    @property
    def n_owned_cells(self):
        return self.part.n_owned_cells if self.part is not None else self.num_cells()


# --------------------------------------------------------------------------------------
# generators (layouts: SURVEY 8d; verified against the reference's tangling KATs and demo traces)
# --------------------------------------------------------------------------------------
def square_arrays(nx, ny, Lx=1.0, Ly=1.0):
    xs = np.linspace(0.0, Lx, nx + 1)
    ys = np.linspace(0.0, Ly, ny + 1)
    X, Y = np.meshgrid(xs, ys, indexing="ij")
    coords = np.stack([X.ravel(), Y.ravel()], axis=1)
    ii, jj = np.meshgrid(np.arange(nx, dtype=np.int64), np.arange(ny, dtype=np.int64), indexing="ij")
    q0 = (ii * (ny + 1) + jj).ravel()
    q1, q3 = q0 + 1, q0 + (ny + 1)
    q2 = q3 + 1
    cells = np.empty((2 * nx * ny, 3), dtype=np.int32)
    cells[0::2, 0], cells[0::2, 1], cells[0::2, 2] = q0, q1, q3
    cells[1::2, 0], cells[1::2, 1], cells[1::2, 2] = q1, q2, q3
    j = np.arange(ny)
    i = np.arange(nx)
    bf = np.concatenate([
        np.stack([j, j + 1], axis=1),
        np.stack([nx * (ny + 1) + j, nx * (ny + 1) + j + 1], axis=1),
        np.stack([i * (ny + 1), (i + 1) * (ny + 1)], axis=1),
        np.stack([i * (ny + 1) + ny, (i + 1) * (ny + 1) + ny], axis=1)]).astype(np.int32)
    bt = np.concatenate([np.full(ny, 1), np.full(ny, 2), np.full(nx, 3), np.full(nx, 4)]).astype(np.int32)
    return coords, cells, bf, bt


_KUHN = np.array([[0, 1, 3, 7], [0, 1, 7, 5], [0, 5, 7, 4], [0, 3, 2, 7], [0, 6, 4, 7], [0, 2, 6, 7]])


def cube_arrays(nx, ny, nz):
    xs, ys, zs = (np.linspace(0.0, 1.0, n + 1) for n in (nx, ny, nz))
    Z, Y, X = np.meshgrid(zs, ys, xs, indexing="ij")
    coords = np.stack([X.ravel(), Y.ravel(), Z.ravel()], axis=1)
    sx, sy, sz = 1, nx + 1, (nx + 1) * (ny + 1)
    kk, jj, ii = np.meshgrid(np.arange(nz, dtype=np.int64), np.arange(ny, dtype=np.int64),
                             np.arange(nx, dtype=np.int64), indexing="ij")
    v0 = (kk * sz + jj * sy + ii).ravel()
    off = np.array([0, sx, sy, sx + sy, sz, sx + sz, sy + sz, sx + sy + sz], dtype=np.int64)
    cells = np.empty((v0.size, 6, 4), dtype=np.int32)
    for t in range(6):
        for l in range(4):
            cells[:, t, l] = v0 + off[_KUHN[t, l]]
    cells = cells.reshape(-1, 4)
    # boundary faces, per bounding plane, straight from the hex faces (two triangles each):
    # the Kuhn split of a hex face (a,b,c,d) with a the lowest and d the highest corner is (a,b,d),(a,c,d)
    def quad_faces(a, b, c, d):
        return np.concatenate([np.stack([a, b, d], axis=1), np.stack([a, c, d], axis=1)])

    def grid(n1, n2):
        p, q = np.meshgrid(np.arange(n1, dtype=np.int64), np.arange(n2, dtype=np.int64), indexing="ij")
        return p.ravel(), q.ravel()

    bf, bt = [], []
    j, k = grid(ny, nz)
    for tag, i0 in ((1, 0), (2, nx)):
        a = k * sz + j * sy + i0
        bf.append(quad_faces(a, a + sy, a + sz, a + sy + sz)); bt.append(np.full(2 * a.size, tag))
    i, k = grid(nx, nz)
    for tag, j0 in ((3, 0), (4, ny)):
        a = k * sz + j0 * sy + i
        bf.append(quad_faces(a, a + sx, a + sz, a + sx + sz)); bt.append(np.full(2 * a.size, tag))
    i, j = grid(nx, ny)
    for tag, k0 in ((5, 0), (6, nz)):
        a = k0 * sz + j * sy + i
        bf.append(quad_faces(a, a + sx, a + sy, a + sx + sy)); bt.append(np.full(2 * a.size, tag))
    return coords, cells, np.concatenate(bf).astype(np.int32), np.concatenate(bt).astype(np.int32)


def cube_part(nx, ny, nz, rank, size):
    """Rank ``rank``'s part of ``cube_arrays(nx, ny, nz)`` under the block partition, generated WITHOUT the global mesh:
    only the hexahedron layers that can touch an owned vertex (vertex numbers run x fastest, z slowest, so an index block is
    a z-slab) are created, in the global order, and cut by the same routine as a global mesh (``partition._cut_block``).
    The result is identical, array by array, to ``partition_mesh(*cube_arrays(nx, ny, nz), rank, size)``
    (tests/test_partition.py); at 256^3 on 8 ranks each rank touches 1/7 of the 100 M cells instead of all of them."""
    from .partition import _cut_block, block_ranges

    xs, ys, zs = (np.linspace(0.0, 1.0, n + 1) for n in (nx, ny, nz))
    sx, sy, sz = 1, nx + 1, (nx + 1) * (ny + 1)
    V = sz * (nz + 1)
    starts = block_ranges(V, size)
    lo, hi = int(starts[rank]), int(starts[rank + 1])
    klo, khi = lo // sz, (max(hi, lo + 1) - 1) // sz          # z-layers of the owned vertices
    k0, k1 = max(klo - 1, 0), min(khi, nz - 1)                 # hexahedron layers that touch them
    kk, jj, ii = np.meshgrid(np.arange(k0, k1 + 1, dtype=np.int64), np.arange(ny, dtype=np.int64),
                             np.arange(nx, dtype=np.int64), indexing="ij")
    v0 = (kk * sz + jj * sy + ii).ravel()
    off = np.array([0, sx, sy, sx + sy, sz, sx + sz, sy + sz, sx + sy + sz], dtype=np.int64)
    cells = np.empty((v0.size, 6, 4), dtype=np.int32)
    for t in range(6):
        for l in range(4):
            cells[:, t, l] = v0 + off[_KUHN[t, l]]
    cells = cells.reshape(-1, 4)
    cell_gid = np.int64(k0) * ny * nx * 6 + np.arange(cells.shape[0], dtype=np.int64)

    def quad_faces(a, b, c, d):
        return np.concatenate([np.stack([a, b, d], axis=1), np.stack([a, c, d], axis=1)])

    def grid(r1, r2):
        p, q = np.meshgrid(r1, r2, indexing="ij")
        return p.ravel(), q.ravel()

    kr = np.arange(k0, k1 + 1, dtype=np.int64)
    bf, bt = [], []
    j, k = grid(np.arange(ny, dtype=np.int64), kr)
    for tag, i0 in ((1, 0), (2, nx)):
        a = k * sz + j * sy + i0
        bf.append(quad_faces(a, a + sy, a + sz, a + sy + sz)); bt.append(np.full(2 * a.size, tag))
    i, k = grid(np.arange(nx, dtype=np.int64), kr)
    for tag, j0 in ((3, 0), (4, ny)):
        a = k * sz + j0 * sy + i
        bf.append(quad_faces(a, a + sx, a + sz, a + sx + sz)); bt.append(np.full(2 * a.size, tag))
    i, j = grid(np.arange(nx, dtype=np.int64), np.arange(ny, dtype=np.int64))
    for tag, kp, wanted in ((5, 0, klo == 0), (6, nz, khi == nz)):
        if wanted:
            a = kp * sz + j * sy + i
            bf.append(quad_faces(a, a + sx, a + sy, a + sx + sy)); bt.append(np.full(2 * a.size, tag))

    def coords_of(gv):
        gv = np.asarray(gv, dtype=np.int64)
        return np.stack([xs[gv % sy], ys[(gv // sy) % (ny + 1)], zs[gv // sz]], axis=1)

    return _cut_block(coords_of, 3, V, cells, cell_gid, 6 * nx * ny * nz, np.concatenate(bf).astype(np.int32),
                      np.concatenate(bt).astype(np.int32), rank, size, starts)


def square_part(nx, ny, rank, size, Lx=1.0, Ly=1.0):
    """The same for ``square_arrays`` (vertex numbers run y fastest: an index block is an x-slab)."""
    from .partition import _cut_block, block_ranges

    xs = np.linspace(0.0, Lx, nx + 1)
    ys = np.linspace(0.0, Ly, ny + 1)
    sy = ny + 1
    V = (nx + 1) * sy
    starts = block_ranges(V, size)
    lo, hi = int(starts[rank]), int(starts[rank + 1])
    ilo, ihi = lo // sy, (max(hi, lo + 1) - 1) // sy
    i0, i1 = max(ilo - 1, 0), min(ihi, nx - 1)
    ii, jj = np.meshgrid(np.arange(i0, i1 + 1, dtype=np.int64), np.arange(ny, dtype=np.int64), indexing="ij")
    q0 = (ii * sy + jj).ravel()
    q1, q3 = q0 + 1, q0 + sy
    q2 = q3 + 1
    cells = np.empty((2 * q0.size, 3), dtype=np.int32)
    cells[0::2, 0], cells[0::2, 1], cells[0::2, 2] = q0, q1, q3
    cells[1::2, 0], cells[1::2, 1], cells[1::2, 2] = q1, q2, q3
    cell_gid = np.int64(i0) * ny * 2 + np.arange(cells.shape[0], dtype=np.int64)
    j = np.arange(ny, dtype=np.int64)
    i = np.arange(i0, i1 + 1, dtype=np.int64)
    bf, bt = [], []
    if ilo == 0:
        bf.append(np.stack([j, j + 1], axis=1)); bt.append(np.full(ny, 1))
    if ihi == nx:
        bf.append(np.stack([nx * sy + j, nx * sy + j + 1], axis=1)); bt.append(np.full(ny, 2))
    bf.append(np.stack([i * sy, (i + 1) * sy], axis=1)); bt.append(np.full(i.size, 3))
    bf.append(np.stack([i * sy + ny, (i + 1) * sy + ny], axis=1)); bt.append(np.full(i.size, 4))

    def coords_of(gv):
        gv = np.asarray(gv, dtype=np.int64)
        return np.stack([xs[gv // sy], ys[gv % sy]], axis=1)

    return _cut_block(coords_of, 2, V, cells, cell_gid, 2 * nx * ny, np.concatenate(bf).astype(np.int32),
                      np.concatenate(bt).astype(np.int32), rank, size, starts)


def _maybe_distribute(arrays, distribute, part_fn=None):
    from . import parallel

    if distribute is None:
        distribute = parallel.is_active()
    if callable(arrays) and not (distribute and part_fn is not None and os.environ.get("MV_PARTITION", "block") == "block"):
        arrays = arrays()
    if not distribute:
        coords, cells, bf, bt = arrays
        return Mesh(coords, cells, bf, bt)
    from .partition import partition_mesh

    if callable(arrays):      # generated mesh, block partition: this rank's slab only, never the global mesh
        lp = part_fn(parallel.rank(), parallel.size())
    else:
        coords, cells, bf, bt = arrays
        lp = partition_mesh(coords, cells, bf, bt, parallel.rank(), parallel.size())
    return Mesh(lp.coords, lp.cells, lp.bfacets, lp.btags, part=lp)


def UnitSquareMesh(nx, ny=None, distribute=None):
    ny = nx if ny is None else ny
    return _maybe_distribute(lambda: square_arrays(nx, ny), distribute, lambda r, s: square_part(nx, ny, r, s))


def RectangleMesh(nx, ny, Lx, Ly, distribute=None):
    return _maybe_distribute(lambda: square_arrays(nx, ny, Lx, Ly), distribute, lambda r, s: square_part(nx, ny, r, s, Lx, Ly))


def UnitCubeMesh(nx, ny=None, nz=None, distribute=None):
    ny = nx if ny is None else ny
    nz = nx if nz is None else nz
    return _maybe_distribute(lambda: cube_arrays(nx, ny, nz), distribute, lambda r, s: cube_part(nx, ny, nz, r, s))


def UnitTriangleMesh():
    return Mesh(np.array([[0.0, 0.0], [1.0, 0.0], [0.0, 1.0]]), np.array([[0, 1, 2]]),
                np.array([[1, 2], [0, 2], [0, 1]]), np.array([1, 1, 1]))


def interval_arrays(n):
    return (np.linspace(0.0, 1.0, n + 1)[:, None], np.stack([np.arange(n), np.arange(n) + 1], axis=1).astype(np.int32),
            np.array([[0], [n]], dtype=np.int32), np.array([1, 2], dtype=np.int32))


def UnitIntervalMesh(n):
    return Mesh(*interval_arrays(n))


def _periodic(arrays, shape):
    """Identify opposite faces of a structured unit mesh numbered on the vertex lattice `shape` (first index slowest):
    periodic P1 cell-node map + DG coordinates (PeriodicUnit*Mesh, both / all directions)."""
    from . import parallel

    if parallel.is_active():
        raise NotImplementedError("periodic meshes are not partitioned over several GPUs yet")
    coords, cells = arrays[0], arrays[1]
    per = np.arange(int(np.prod([s - 1 for s in shape]))).reshape([s - 1 for s in shape])
    for ax in range(len(shape)):
        per = np.concatenate([per, np.take(per, [0], axis=ax)], axis=ax)
    pcells = per.ravel()[cells].astype(np.int32)
    return Mesh(coords[cells].reshape(-1, coords.shape[1]), pcells, None, None, periodic=True)


def PeriodicUnitIntervalMesh(n):
    return _periodic(interval_arrays(n), (n + 1,))


def PeriodicUnitSquareMesh(nx, ny=None):
    """demos/monge_ampere_periodic.py:16."""
    ny = nx if ny is None else ny
    return _periodic(square_arrays(nx, ny), (nx + 1, ny + 1))


def PeriodicUnitCubeMesh(nx, ny=None, nz=None):
    ny = nx if ny is None else ny
    nz = nx if nz is None else nz
    return _periodic(cube_arrays(nx, ny, nz), (nz + 1, ny + 1, nx + 1))


def from_firedrake(fmesh):
    """Ingest a live Firedrake mesh (P1 coordinates only; SURVEY 8b 'what is ingested')."""
    coords = np.array(fmesh.coordinates.dat.data_ro_with_halos, dtype=np.float64)
    cells = np.array(fmesh.coordinates.cell_node_map().values_with_halo, dtype=np.int32)
    ef = fmesh.exterior_facets
    d = coords.shape[1] if coords.ndim > 1 else 1
    # facet -> vertices: the facet's cell minus the vertex opposite the local facet number
    fcells = np.asarray(ef.facet_cell).reshape(-1)
    lfn = np.asarray(ef.local_facet_dat.data_ro).reshape(-1)
    bf = np.stack([np.delete(cells[c], l) for c, l in zip(fcells, lfn)]) if len(fcells) else np.zeros((0, d), np.int32)
    bt = np.asarray(ef.markers, dtype=np.int32) if ef.markers is not None else np.zeros(len(fcells), np.int32)
    deg = fmesh.coordinates.ufl_element().degree()
    return Mesh(coords, cells, bf, bt, degree=deg)
