"""Thin object layer over the C ABI: device-resident mesh operators and mover state.

PyTorch is used for tensor ownership (user-visible fields), streams and
``torch.distributed``; every numerical step is a call into libmovement_b200.so.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib as L


def _require_cuda():
    if not torch.cuda.is_available():
        raise RuntimeError("movement_b200 needs a CUDA device (sm_100a); there is no CPU fallback")


def use_current_stream():
    """Enqueue subsequent library work on torch's current CUDA stream."""
    L.check(L.load().mv_set_stream(C.c_void_p(torch.cuda.current_stream().cuda_stream)))


class DeviceMesh:
    """Ingested mesh: topology and the constant operators on the computational mesh.

    ``coords`` [V,d] f64, ``cells`` [C,d+1] i32, ``bfacets`` [F,d] i32, ``btags`` [F] i32 may
    be numpy arrays (host) or CUDA tensors.
    """

    def __init__(self, coords, cells, bfacets=None, btags=None, n_owned_vertices=None, n_owned_cells=None, periodic=False,
                 num_vertices=None):
        _require_cuda()
        lib = L.load()
        self.dim = int(coords.shape[1])
        self.periodic = bool(periodic)
        # periodic: coords is the DG1 coordinate field [C*(d+1), d], cells the periodic P1 map over num_vertices vertices
        self.V = int(num_vertices) if periodic else int(coords.shape[0])
        self.C = int(cells.shape[0])
        self.F = 0 if bfacets is None else int(bfacets.shape[0])
        self.Vown = self.V if n_owned_vertices is None else int(n_owned_vertices)
        self.Cown = self.C if n_owned_cells is None else int(n_owned_cells)
        self.device = torch.device("cuda", torch.cuda.current_device())
        L.check(lib.mv_set_device(torch.cuda.current_device()))
        use_current_stream()
        coords = self._prep(coords, np.float64, torch.float64)
        cells = self._prep(cells, np.int32, torch.int32)
        bfacets = self._prep(bfacets, np.int32, torch.int32)
        btags = self._prep(btags, np.int32, torch.int32)
        h = C.c_void_p()
        L.check(lib.mv_mesh_create_ex(self.dim, self.V, self.C, L.ptr(coords), L.ptr(cells), self.F, L.ptr(bfacets),
                                      L.ptr(btags), self.Vown, self.Cown, 1 if self.periodic else 0, C.byref(h)))
        self.handle = h
        info = (C.c_int64 * 16)()
        L.check(lib.mv_mesh_info(h, info))
        self.nnz, self.sell_len, self.nslices = int(info[3]), int(info[4]), int(info[5])
        sc = (C.c_double * 3)()
        L.check(lib.mv_mesh_scalars(h, sc))
        self.total_volume, self.ml_sumsq, self.ml_sum = float(sc[0]), float(sc[1]), float(sc[2])

    @staticmethod
    def _prep(a, npdt, tdt):
        if a is None:
            return None
        if isinstance(a, torch.Tensor):
            return a.to(tdt).contiguous()
        return np.ascontiguousarray(a, dtype=npdt)

    def __del__(self):
        h = getattr(self, "handle", None)
        if h is not None and L is not None and L._lib is not None:
            L._lib.mv_mesh_destroy(h)
            self.handle = None

    # ---- helpers ------------------------------------------------------------------
    def empty(self, *shape, dtype=torch.float64):
        return torch.empty(*shape, dtype=dtype, device=self.device)

    def zeros(self, *shape, dtype=torch.float64):
        return torch.zeros(*shape, dtype=dtype, device=self.device)

    def get(self, what):
        """Copy an internal array to the host (tests / inspection)."""
        d = self.dim
        spec = {
            L.MV_GET_ROW_NNZ: (np.int32, (self.Vown,)),
            L.MV_GET_SLICE_PTR: (np.int64, (self.nslices + 1,)),
            L.MV_GET_COLS: (np.int32, (self.sell_len,)),
            L.MV_GET_MASS: (np.float64, (self.sell_len,)),
            L.MV_GET_STIFF: (np.float64, (self.sell_len,)),
            L.MV_GET_LUMPED: (np.float64, (self.Vown,)),
            L.MV_GET_CELL_VOL: (np.float64, (self.C,)),
            L.MV_GET_BF_NORMAL: (np.float64, (self.F, d)),
            L.MV_GET_BF_AREA: (np.float64, (self.F,)),
            L.MV_GET_BF_CELL: (np.int32, (self.F,)),
        }[what]
        out = np.empty(spec[1], dtype=spec[0])
        L.check(L.load().mv_mesh_get(self.handle, what, L.ptr(out)))
        return out

    def to_scipy(self, what):
        """Rebuild the assembled operator (MV_GET_MASS / MV_GET_STIFF) as scipy CSR."""
        import scipy.sparse as sp

        sptr = self.get(L.MV_GET_SLICE_PTR)
        cols = self.get(L.MV_GET_COLS)
        vals = self.get(what)
        nnz = self.get(L.MV_GET_ROW_NNZ)
        rows_i, cols_i, vals_i = [], [], []
        rows = np.arange(self.Vown)
        s = rows // 32
        lane = rows % 32
        base = sptr[s] + lane
        for k in range(int(nnz.max())):
            sel = nnz > k
            pos = base[sel] + k * 32
            rows_i.append(rows[sel]); cols_i.append(cols[pos]); vals_i.append(vals[pos])
        return sp.csr_matrix((np.concatenate(vals_i), (np.concatenate(rows_i), np.concatenate(cols_i))),
                             shape=(self.Vown, self.V))

    # ---- linear algebra -------------------------------------------------------------
    def spmv(self, op, x):
        ncomp = 1 if x.dim() == 1 else int(np.prod(x.shape[1:]))
        y = self.empty((self.Vown,) + tuple(x.shape[1:]))
        use_current_stream()
        L.check(L.load().mv_spmv(self.handle, op, ncomp, L.ptr(x), L.ptr(y)))
        return y

    def pcg(self, op, rhs, x, mask=None, rtol=1e-12, atol=1e-50, maxit=10000, project_mean=False):
        """Solve in place; returns (iterations, per-component relative residuals)."""
        ncomp = 1 if x.dim() == 1 else int(np.prod(x.shape[1:]))
        it = C.c_int(0)
        rel = (C.c_double * ncomp)()
        use_current_stream()
        L.check(L.load().mv_pcg_solve(self.handle, op, ncomp, L.ptr(rhs), L.ptr(x), L.ptr(mask), rtol, atol, maxit,
                                      1 if project_mean else 0, C.byref(it), rel))
        return it.value, np.array(rel[:])

    def slice_schedule(self):
        """State of the slice schedule of the gather kernels (csrc/mv_sched.cu): "none" before the first
        mass solve, "identity", or "tiled" (lattice tiles of 16 neighbouring slices)."""
        info = (C.c_int64 * 16)()
        L.check(L.load().mv_mesh_info(self.handle, info))
        return ("none", "identity", "tiled")[int(info[10])]

    def set_preconditioner(self, kind):
        """"amg" (default) or "jacobi" for the stiffness-type solves (mass solves are always Jacobi)."""
        k = {"amg": L.MV_PC_AMG, "jacobi": L.MV_PC_JACOBI}.get(kind, kind)
        L.check(L.load().mv_set_preconditioner(self.handle, int(k)))

    def amg_info(self, op=L.MV_OP_STIFFNESS):
        """Hierarchy of the cached multigrid preconditioner of `op` (None before its first solve)."""
        info = (C.c_int64 * 40)()
        rho = (C.c_double * 16)()
        ms = C.c_double(0.0)
        L.check(L.load().mv_amg_info(self.handle, op, info, 16, rho, C.byref(ms)))
        nl = int(info[0])
        if nl == 0:
            return None
        return {"levels": nl, "coarsest_rows": int(info[1]), "dense_coarse_solve": bool(info[2]),
                "rows": [int(info[3 + 2 * l]) for l in range(nl)], "nnz": [int(info[4 + 2 * l]) for l in range(nl)],
                "rho": [float(rho[l]) for l in range(nl)], "setup_ms": float(ms.value)}

    def amg_matrix(self, level, which="A", op=L.MV_OP_STIFFNESS, ncols=None):
        """One matrix of the cached hierarchy as scipy CSR (tests / inspection): "A", "P" or "R"."""
        import scipy.sparse as sp

        w = {"A": 0, "P": 1, "R": 2}[which]
        size = (C.c_int64 * 4)()
        lib = L.load()
        L.check(lib.mv_amg_level_get(self.handle, op, level, w, size, None, None, None, None))
        nrows, nslices, ln = int(size[0]), int(size[1]), int(size[2])
        sptr = np.empty(nslices + 1, np.int64); cols = np.empty(ln, np.int32)
        vals = np.empty(ln, np.float64); nnz = np.empty(nrows, np.int32)
        L.check(lib.mv_amg_level_get(self.handle, op, level, w, size, L.ptr(sptr), L.ptr(cols), L.ptr(vals), L.ptr(nnz)))
        rows = np.arange(nrows)
        base = sptr[rows // 32] + rows % 32
        ri, ci, vi = [], [], []
        for k in range(int(nnz.max()) if nrows else 0):
            sel = nnz > k
            pos = base[sel] + k * 32
            ri.append(rows[sel]); ci.append(cols[pos]); vi.append(vals[pos])
        if ncols is None:
            ncols = int(max(c.max() for c in ci)) + 1
        return sp.csr_matrix((np.concatenate(vi), (np.concatenate(ri), np.concatenate(ci))), shape=(nrows, ncols))

    # ---- tangling ---------------------------------------------------------------------
    def tangle_init(self, coords):
        detJ0 = self.empty(self.C)
        use_current_stream()
        L.check(L.load().mv_tangle_init(self.handle, L.ptr(coords), L.ptr(detJ0)))
        return detJ0

    def tangle_check(self, coords, detJ0, ratio=None):
        n = C.c_int64(0)
        use_current_stream()
        L.check(L.load().mv_tangle_check(self.handle, L.ptr(coords), L.ptr(detJ0), L.ptr(ratio), C.byref(n)))
        return int(n.value)

    def cell_volumes(self, coords, volume):
        st = (C.c_double * 4)()
        use_current_stream()
        L.check(L.load().mv_cell_volumes(self.handle, L.ptr(coords), L.ptr(volume), st))
        return tuple(st[:])  # max, min, mean, cov


class MAState:
    """Monge-Ampere mover state: torch owns the fields, the library owns the operators."""

    FIELDS = {
        "xi": (L.MV_F_XI, "vec"), "x": (L.MV_F_X, "vec"), "phi": (L.MV_F_PHI, "scal"), "phi_old": (L.MV_F_PHI_OLD, "scal"),
        "H": (L.MV_F_H, "ten"), "H_old": (L.MV_F_H_OLD, "ten"), "grad_phi": (L.MV_F_GRAD_PHI, "vec"),
        "monitor": (L.MV_F_MONITOR, "scal"), "volume": (L.MV_F_VOLUME, "cell"), "original_volume": (L.MV_F_ORIG_VOLUME, "cell"),
        "residual": (L.MV_F_RESIDUAL, "scal"),
    }

    def __init__(self, dmesh: DeviceMesh, coords):
        self.dmesh = dmesh
        lib = L.load()
        h = C.c_void_p()
        use_current_stream()
        L.check(lib.mv_ma_create(dmesh.handle, C.byref(h)))
        self.handle = h
        V, Cc, d = dmesh.V, dmesh.C, dmesh.dim
        shapes = {"vec": (V, d), "scal": (V,), "ten": (V, d, d), "cell": (Cc,)}
        for name, (fid, kind) in self.FIELDS.items():
            t = dmesh.zeros(shapes[kind])
            setattr(self, name, t)
            L.check(lib.mv_ma_bind(h, fid, L.ptr(t)))
        xi = torch.as_tensor(np.ascontiguousarray(coords, dtype=np.float64)) if not isinstance(coords, torch.Tensor) else coords
        self.xi.copy_(xi.to(dmesh.device))
        self.x.copy_(self.xi)
        self.original_volume.copy_(torch.as_tensor(dmesh.get(L.MV_GET_CELL_VOL)).to(dmesh.device))
        self.volume.copy_(self.original_volume)
        self.monitor.fill_(1.0)

    def __del__(self):
        h = getattr(self, "handle", None)
        if h is not None and L is not None and L._lib is not None:
            L._lib.mv_ma_destroy(h)
            self.handle = None

    def set_boundary(self, fixed_tags=()):
        tags = np.ascontiguousarray(list(fixed_tags), dtype=np.int32)
        use_current_stream()
        L.check(L.load().mv_ma_set_boundary(self.handle, len(tags), L.ptr(tags) if len(tags) else None))

    def bc_mask(self):
        out = np.empty(self.dmesh.Vown, dtype=np.uint16)
        L.check(L.load().mv_ma_get_bcmask(self.handle, L.ptr(out)))
        return out

    def set_solver(self, rtol=1e-12, atol=1e-50, maxit_mass=1000, maxit_poisson=200000):
        L.check(L.load().mv_ma_set_solver(self.handle, rtol, atol, maxit_mass, maxit_poisson))

    def set_monitor(self, kind, params=()):
        arr = (C.c_double * max(len(params), 1))(*params)
        L.check(L.load().mv_ma_set_monitor(self.handle, kind, len(params), arr))

    def l2_project(self):
        it = C.c_int(0)
        use_current_stream()
        L.check(L.load().mv_ma_l2_project(self.handle, C.byref(it)))
        return it.value

    def update_coordinates(self):
        use_current_stream()
        L.check(L.load().mv_ma_update_coordinates(self.handle))

    def eval_monitor(self, coords, out):
        use_current_stream()
        L.check(L.load().mv_ma_eval_monitor(self.handle, L.ptr(coords), L.ptr(out)))

    def diagnostics(self, use_H_old=True):
        out = (C.c_double * 6)()
        use_current_stream()
        L.check(L.load().mv_ma_diagnostics(self.handle, 1 if use_H_old else 0, out))
        return dict(theta=out[0], residual=out[1], vol_max=out[2], vol_min=out[3], vol_mean=out[4], cov=out[5])

    def pseudotimestep(self, dt):
        it = C.c_int(0)
        use_current_stream()
        L.check(L.load().mv_ma_pseudotimestep(self.handle, dt, C.byref(it)))
        return it.value

    def equidistribute(self):
        it = C.c_int(0)
        use_current_stream()
        L.check(L.load().mv_ma_equidistribute(self.handle, C.byref(it)))
        return it.value

    def advance(self):
        use_current_stream()
        L.check(L.load().mv_ma_advance(self.handle))

    def relax_iterate(self, dt, niter):
        trace = (C.c_double * (6 * niter))()
        its = (C.c_int * (3 * niter))()
        use_current_stream()
        L.check(L.load().mv_ma_relax_iterate(self.handle, dt, niter, trace, its))
        return np.array(trace[:]).reshape(niter, 6), np.array(its[:]).reshape(niter, 3)


def device_mesh_from(mesh):
    """Ingest a :class:`movement_b200.mesh.Mesh` (one-off)."""
    part = mesh.part
    dm = DeviceMesh(mesh.coordinates.tensor, mesh.cells, mesh.bfacets if len(mesh.bfacets) else None,
                    mesh.btags if len(mesh.btags) else None,
                    n_owned_vertices=mesh.n_owned_vertices, n_owned_cells=mesh.n_owned_cells,
                    periodic=getattr(mesh, "periodic", False), num_vertices=mesh.num_vertices())
    if part is not None:
        lib = L.load()
        nn = len(part.neighbours)
        L.check(lib.mv_mesh_set_halo(dm.handle, nn, L.ptr(np.ascontiguousarray(part.neighbours, dtype=np.int32)),
                                     L.ptr(np.ascontiguousarray(part.send_ptr, dtype=np.int64)),
                                     L.ptr(np.ascontiguousarray(part.send_idx, dtype=np.int32)),
                                     L.ptr(np.ascontiguousarray(part.recv_ptr, dtype=np.int64))))
        L.check(lib.mv_mesh_set_cell_order(dm.handle, L.ptr(np.ascontiguousarray(part.global_cell, dtype=np.int64))))
        dm.global_cells = part.global_num_cells
        dm.global_vertices = part.global_num_vertices
    else:
        dm.global_cells, dm.global_vertices = dm.C, dm.V
    return dm
