"""LaplacianSmoother -- drop-in for movement/laplacian_smoothing.py:13-127."""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib as L
from .core import use_current_stream
from .mesh import DirichletBC, Function
from .mover import PrimeMover, _print
from .profiling import event

__all__ = ["LaplacianSmoother"]


def _as_bc_list(bcs):
    if isinstance(bcs, DirichletBC):
        return [bcs]
    return list(bcs)


def dirichlet_mask_and_values(bcs, dim, n_vertices, device, out=None):
    """uint16 component mask per vertex + [V,d] boundary values; later BCs override earlier ones at
    shared vertices (movement/spring.py:225-230; Firedrake applies BCs in order as well)."""
    mask = np.zeros(n_vertices, dtype=np.uint16)
    vals = torch.zeros((n_vertices, dim), dtype=torch.float64, device=device) if out is None else out
    for bc in bcs:
        nodes = torch.as_tensor(bc.nodes, device=device)
        mask[bc.nodes] = (1 << dim) - 1
        if len(bc.nodes):
            vals[nodes] = bc.values()[nodes]
    return mask, vals


class LaplacianSmoother(PrimeMover):
    r"""Vector Laplace mesh velocity, :math:`-\Delta v = 0` with Dirichlet data, then
    :math:`x \leftarrow x + v\,\Delta t` (movement/laplacian_smoothing.py:13-34)."""

    # Firedrake's default for the cached LinearVariationalSolver is ksp_rtol 1e-7;
    # the stiffness matrix lives on the computational mesh and never changes.
    ksp_rtol = 1.0e-12
    ksp_max_it = 100000

    @event("movement.laplacian_smoothing.LaplacianSmoother.__init__")
    def __init__(self, mesh, timestep, **kwargs):
        super().__init__(mesh, **kwargs)
        assert timestep > 0.0
        self.dt = timestep
        self.displacement = np.zeros((self.mesh.num_vertices(), self.dim))
        self.cg_iterations = []

    def _create_functions(self):
        super()._create_functions()
        self.rhs = Function(self.coord_space, name="Zero RHS")

    def _solve(self, boundary_conditions, solver_parameters=None):
        """K v = 0 with the Dirichlet rows of ``boundary_conditions``.  Like the reference's cached
        solver (laplacian_smoothing.py:64-77) the BC *objects* and solver parameters of the first
        call are kept; their Functions are shared by reference, so updated values are seen."""
        if not hasattr(self, "_solver"):
            sp = dict(solver_parameters or {})
            self._solver = {
                "bcs": _as_bc_list(boundary_conditions),
                "rtol": float(sp.get("ksp_rtol", self.ksp_rtol)),
                "atol": float(sp.get("ksp_atol", 1e-300)),
                "maxit": int(sp.get("ksp_max_it", self.ksp_max_it)),
            }
            for bc in self._solver["bcs"]:
                if bc.function_space() != self.coord_space:
                    raise ValueError(f"Boundary conditions must have {type(self).__name__}.coord_space"
                                     " as their function space.")
        s = self._solver
        V = self.mesh.num_vertices()
        vt = self.v.tensor
        vt.zero_()
        mask, _ = dirichlet_mask_and_values(s["bcs"], self.dim, V, self.mesh.device, out=vt)
        self._mask_dev = torch.as_tensor(mask.view(np.int16), device=self.mesh.device)
        it = C.c_int(0)
        nrm = C.c_double(0.0)
        use_current_stream()
        st = L.load().mv_laplace_move(self._dm.handle, L.ptr(vt), L.ptr(self._mask_dev), float(self.dt),
                                      L.ptr(self.x.tensor), s["rtol"], s["atol"], s["maxit"], C.byref(it), C.byref(nrm))
        self.cg_iterations.append(it.value)
        L.check(st)
        self.v.written()
        self.x.written()
        self._displacement_norm = nrm.value

    @event("movement.laplacian_smoothing.LaplacianSmoother.move")
    def move(self, time, update_boundary_velocity=None, boundary_conditions=None, solver_parameters=None):
        """movement/laplacian_smoothing.py:80-127."""
        if update_boundary_velocity is not None:
            update_boundary_velocity(time)
        if not boundary_conditions:
            boundary_conditions = DirichletBC(self.coord_space, 0, "on_boundary")

        self.mesh.coordinates.assign(self.xi)
        try:
            self._solve(boundary_conditions, solver_parameters=solver_parameters)
        except L.MovementB200Error as conv_err:
            if conv_err.status != L.MV_ERR_NOT_CONVERGED:
                raise
            self._convergence_error(exception=conv_err)

        self.mesh.coordinates.assign(self.x)
        self._volume_stats = self._dm.cell_volumes(self.mesh.coordinates.tensor, self.volume.tensor)
        self.volume.written()
        _print(f"{time:.2f} s"
               f"   Volume ratio {self.volume_ratio:5.2f}"
               f"   Variation (σ/μ) {self.coefficient_of_variation:8.2e}"
               f"   Displacement {self._displacement_norm:.2f} m")
        if hasattr(self, "tangling_checker"):
            self.tangling_checker.check()

    @property
    def displacement(self):
        """v * dt as a host array (the reference keeps a numpy array, laplacian_smoothing.py:116)."""
        if hasattr(self, "v") and getattr(self, "_displacement_norm", None) is not None:
            return self.v.dat.data_ro * self.dt
        return self._displacement0

    @displacement.setter
    def displacement(self, value):
        self._displacement0 = value
