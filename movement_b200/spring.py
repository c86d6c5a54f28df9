"""SpringMover -- drop-in for movement/spring.py:15-298 (lineal springs).

The reference assembles a *dense* 2V x 2V numpy matrix in a Python edge loop and calls LAPACK
(spring.py:148-181,266): O(V^2) memory.  Here the same row-replaced system is applied
matrix-free from the edge geometry (one kernel per CG iteration) and solved with a
block-Jacobi PCG on the device, so the 4096 x 4096 configuration fits one GPU.
"""
from __future__ import annotations

import ctypes as C
from collections.abc import Iterable
from types import SimpleNamespace

import numpy as np
import torch

from . import _lib as L
from .core import use_current_stream
from .mesh import DirichletBC, Function
from .mover import PrimeMover, _print
from .profiling import event

__all__ = ["SpringMover_Lineal", "SpringMover_Torsional", "SpringMover"]


def SpringMover(*args, method="lineal", **kwargs):
    """Factory (movement/spring.py:15-33)."""
    if method == "lineal":
        return SpringMover_Lineal(*args, **kwargs)
    elif method == "torsional":
        return SpringMover_Torsional(*args, **kwargs)
    else:
        raise ValueError(f"Method '{method}' not recognised.")


class _EdgeField:
    """Host view of a per-edge quantity, shaped like ``Function.dat.data`` in the reference."""

    def __init__(self, data):
        self.dat = SimpleNamespace(data=data, data_with_halos=data, data_ro=data)


class SpringMover_Base(PrimeMover):
    """movement/spring.py:36-235."""

    ksp_rtol = 1.0e-12
    ksp_max_it = 200000
    dense_limit = 4096  # vertices; the dense matrix exists only for the reference's small-mesh tests

    def __init__(self, mesh, timestep, **kwargs):
        if mesh.topological_dimension != 2 or mesh.cells.shape[1] != 3:
            raise ValueError("Spring-based Movers only currently support triangular meshes.")
        super().__init__(mesh)
        assert timestep > 0.0
        self.dt = timestep
        num_vertices = self.mesh.num_vertices()
        self._forcing = torch.zeros((num_vertices, 2), dtype=torch.float64, device=self.mesh.device)
        self._mask = np.zeros(num_vertices, dtype=np.uint16)
        self._disp = Function(self.coord_space, name="Displacement")
        self.cg_iterations = []

    @property
    def displacement(self):
        """u * dt, flattened like the reference's numpy vector (spring.py:266); read back lazily."""
        return self._disp.dat.data_ro.reshape(-1)

    # ---- edge quantities (spring.py:70-146) -------------------------------------------------
    def _edge_geometry(self):
        lib = L.load()
        n = C.c_int64(0)
        L.check(lib.mv_spring_edge_count(self._dm.handle, C.byref(n)))
        E = n.value
        edges = np.empty((E, 2), dtype=np.int32)
        length = np.empty(E)
        tangent = np.empty((E, 2))
        angle = np.empty(E)
        use_current_stream()
        L.check(lib.mv_spring_edge_geometry(self._dm.handle, L.ptr(self.mesh.coordinates.tensor), L.ptr(edges),
                                            L.ptr(length), L.ptr(tangent), L.ptr(angle)))
        return edges, length, tangent, angle

    @property
    def edges(self):
        return self._edge_geometry()[0]

    @property
    @event("movement.spring.SpringMover_Base.facet_areas")
    def facet_areas(self):
        return _EdgeField(self._edge_geometry()[1])

    @property
    @event("movement.spring.SpringMover_Base.tangents")
    def tangents(self):
        return _EdgeField(self._edge_geometry()[2])

    @property
    @event("movement.spring.SpringMover_Base.angles")
    def angles(self):
        return _EdgeField(self._edge_geometry()[3])

    @event("movement.spring.SpringMover_Base._stiffness_matrix")
    def _stiffness_matrix(self):
        """Dense 2V x 2V matrix exactly as spring.py:148-181 builds it -- for inspection on small
        meshes only; ``move`` never forms it."""
        Nv = self.mesh.num_vertices()
        if Nv > self.dense_limit:
            raise MemoryError(f"dense stiffness matrix refused for {Nv} vertices; move() is matrix-free")
        edges, length, _, angle = self._edge_geometry()
        K = np.zeros((2 * Nv, 2 * Nv))
        c, s = np.cos(angle), np.sin(angle)
        c2, sc, s2 = c * c / length, s * c / length, s * s / length
        for (i, j), a, b, d in zip(edges, c2, sc, s2):
            blk = np.array([[a, b], [b, d]])
            K[2 * i:2 * i + 2, 2 * i:2 * i + 2] += blk
            K[2 * j:2 * j + 2, 2 * j:2 * j + 2] += blk
            K[2 * i:2 * i + 2, 2 * j:2 * j + 2] -= blk
            K[2 * j:2 * j + 2, 2 * i:2 * i + 2] -= blk
        return K

    def _apply_boundary_conditions(self, boundary_conditions):
        """Validation and forcing/mask update of spring.py:193-234 (without the dense matrix)."""
        boundary_conditions = boundary_conditions or DirichletBC(self.coord_space, 0, "on_boundary")
        if isinstance(boundary_conditions, DirichletBC):
            boundary_conditions = [boundary_conditions]
        assert isinstance(boundary_conditions, Iterable)
        markers = self._all_boundary_segments
        # the reference rebuilds K on every call (spring.py:201), so only THIS call's conditions replace rows;
        # the forcing persists (spring.py:229-230)
        self._mask[:] = 0
        for bc in boundary_conditions:
            if bc.function_space() != self.coord_space:
                raise ValueError(f"Boundary conditions must have {type(self).__name__}.coord_space"
                                 " as their function space.")
            tags = bc.sub_domain
            if tags == ("on_boundary",):
                tags = tuple(markers)
            if not set(tags).issubset(set(markers)):
                raise ValueError(f"{tags} contains invalid boundary tags.")
            nodes = torch.as_tensor(bc.nodes, device=self.mesh.device)
            if len(bc.nodes):
                self._forcing[nodes] = bc.values()[nodes]   # persists between calls (spring.py:229-230)
                self._mask[bc.nodes] = 3
        return boundary_conditions

    @event("movement.spring.SpringMover_Base.assemble_stiffness_matrix")
    def assemble_stiffness_matrix(self, boundary_conditions=None):
        """Dense row-replaced matrix (spring.py:183-235); small meshes only."""
        self._apply_boundary_conditions(boundary_conditions)
        K = self._stiffness_matrix()
        for v in np.nonzero(self._mask)[0]:
            for k in (2 * v, 2 * v + 1):
                K[k][:] = 0
                K[k][k] = 1
        return K


class SpringMover_Lineal(SpringMover_Base):
    """movement/spring.py:238-283."""

    @event("movement.spring.SpringMover_Lineal.move")
    def move(self, time, update_boundary_displacement=None, boundary_conditions=None):
        if update_boundary_displacement is not None:
            update_boundary_displacement(time)
        self._apply_boundary_conditions(boundary_conditions)
        mask_dev = torch.as_tensor(self._mask.view(np.int16), device=self.mesh.device)
        it = C.c_int(0)
        nrm = C.c_double(0.0)
        use_current_stream()
        try:
            st = L.load().mv_spring_move(self._dm.handle, L.ptr(self.mesh.coordinates.tensor), L.ptr(self._forcing),
                                         L.ptr(mask_dev), float(self.dt), L.ptr(self._disp.tensor), self.ksp_rtol, 1e-300,
                                         self.ksp_max_it, C.byref(it), C.byref(nrm))
            self.cg_iterations.append(it.value)
            L.check(st)
        except L.MovementB200Error as conv_err:
            if conv_err.status != L.MV_ERR_NOT_CONVERGED:
                raise
            self._convergence_error(exception=conv_err)
        self.mesh.coordinates.written()
        self._disp.written()
        self._update_plex_coordinates()
        self._volume_stats = self._dm.cell_volumes(self.mesh.coordinates.tensor, self.volume.tensor)
        self.volume.written()
        _print(f"{time:.2f}"
               f"   Volume ratio {self.volume_ratio:5.2f}"
               f"   Variation (σ/μ) {self.coefficient_of_variation:8.2e}"
               f"   Displacement {nrm.value:.2f} m")
        if hasattr(self, "tangling_checker"):
            self.tangling_checker.check()


class SpringMover_Torsional(SpringMover_Lineal):
    """movement/spring.py:286-298."""

    def __init__(self, *args, **kwargs):
        super().__init__(*args, **kwargs)
        raise NotImplementedError("Torsional springs not yet implemented.")
