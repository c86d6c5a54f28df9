"""PrimeMover -- drop-in for movement/mover.py:21-246.

Same constructor arguments, attributes (``mesh``, ``x``, ``xi``, ``v``, ``volume``,
``coord_space``, ``P0``, ``tangling_checker``) and error policy; the DMPlex plumbing
(mover.py:62-75,164-197) is replaced by one ingest into device-resident buffers.
"""
from __future__ import annotations

import abc
from warnings import warn

import numpy as np
import torch

from . import core
from .exceptions import ConvergenceError
from .mesh import Function, Mesh
from .tangling import MeshTanglingChecker

__all__ = ["PrimeMover"]


def _print(*args):
    """Rank-0 print (the reference uses PETSc.Sys.Print)."""
    from . import parallel

    if parallel.rank() == 0:
        print(*args, flush=True)


class PrimeMover(abc.ABC):
    """Base class for all mesh movers."""

    def __init__(self, mesh, monitor_function=None, raise_convergence_errors=True, tangling_check=True,
                 quadrature_degree=None):
        if not torch.cuda.is_available():
            raise RuntimeError("movement_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
        # the user's mesh is never mutated (movement/mover.py:49)
        self.mesh = Mesh(mesh.coordinates.copy(deepcopy=True))
        self.monitor_function = monitor_function
        if not raise_convergence_errors:
            warn(f"{type(self)}.move called with raise_convergence_errors=False."
                 " Beware: this option can produce poor quality meshes!", stacklevel=1)
        self.raise_convergence_errors = raise_convergence_errors
        self.dim = self.mesh.topological_dimension
        self.gdim = self.mesh.geometric_dimension
        # closed-form P1 integrals are exact; a user quadrature degree cannot under-integrate here
        self.quadrature_degree = quadrature_degree

        # ingest once: topology -> device CSR/SELL buffers, M, K, m_l, facet geometry
        self._dm = core.device_mesh_from(self.mesh)

        self._create_function_spaces()
        self._create_functions()
        self._all_boundary_segments = self._global_boundary_tags()

        if tangling_check:
            self.tangling_checker = MeshTanglingChecker(self.mesh, raise_error=raise_convergence_errors,
                                                        _device_mesh=self._dm)

    def _global_boundary_tags(self):
        tags = self.mesh.exterior_facets.unique_markers
        if self.mesh.part is not None:
            import torch.distributed as dist

            out = [None] * dist.get_world_size()
            dist.all_gather_object(out, tags)
            tags = sorted(set(t for ts in out for t in ts))
        return tags

    def _create_function_spaces(self):
        self.coord_space = self.mesh.coordinates.function_space()
        self.P0 = self.mesh.function_space("P0")

    def _create_functions(self):
        self.x = Function(self.mesh.coordinates, name="Physical coordinates")
        self.xi = Function(self.mesh.coordinates, name="Computational coordinates")
        self.v = Function(self.coord_space, name="Mesh velocity")
        self.volume = Function(self.P0, name="Mesh volume")
        self._volume_stats = self._dm.cell_volumes(self.mesh.coordinates.tensor, self.volume.tensor)
        self.volume.written()

    # -- messages / error policy (movement/mover.py:104-162) ------------------------------
    def _convergence_message(self, iterations=None):
        msg = "Solver converged"
        if iterations:
            msg += f" in {iterations} iteration{plural(iterations)}"
        _print(f"{msg}.")

    def _exception(self, msg, exception=None, error_type=ConvergenceError):
        exc_type = error_type if self.raise_convergence_errors else Warning
        if exception:
            raise exc_type(msg) from exception
        raise exc_type(msg)

    def _convergence_error(self, iterations=None, exception=None):
        msg = "Solver failed to converge"
        if iterations:
            msg += f" in {iterations} iteration{plural(iterations)}"
        self._exception(f"{msg}.", exception=exception)

    def _divergence_error(self, iterations=None, exception=None):
        msg = "Solver diverged"
        if iterations:
            msg += f" after {iterations} iteration{plural(iterations)}"
        self._exception(f"{msg}.", exception=exception)

    def _update_plex_coordinates(self):
        """No DMPlex here: the device coordinate buffer *is* the mesh (kept for API parity)."""
        return None

    # -- diagnostics (movement/mover.py:199-217) ------------------------------------------
    # ``_volume_stats`` = (max, min, mean, cov) of the P0 ``volume`` field, produced by the same
    # kernel pass that writes the field (global over ranks).
    @property
    def volume_ratio(self):
        return self._volume_stats[0] / self._volume_stats[1]

    @property
    def coefficient_of_variation(self):
        return self._volume_stats[3]

    @abc.abstractmethod
    def move(self):
        pass  # pragma: no cover

    def to_physical_coordinates(self):
        self.mesh.coordinates.assign(self.x)

    def to_computational_coordinates(self):
        self.mesh.coordinates.assign(self.xi)


def plural(iterations):
    return "s" if iterations != 1 else ""
