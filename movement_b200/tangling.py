"""MeshTanglingChecker -- drop-in for movement/tangling.py:10-53 over the detJ kernel."""
from __future__ import annotations

import warnings

import torch

from .core import device_mesh_from
from .mesh import Function
from .profiling import event

__all__ = ["MeshTanglingChecker"]


class MeshTanglingChecker:
    """Tracks whether any element of ``mesh`` has inverted since the checker was created.

    The check reads whatever ``mesh.coordinates`` currently holds (movement/tangling.py:26-29,37),
    which inside the Monge-Ampere movers is the computational mesh.
    """

    def __init__(self, mesh, raise_error=True, _device_mesh=None):
        self.mesh = mesh
        self.raise_error = raise_error
        self._dm = _device_mesh if _device_mesh is not None else device_mesh_from(mesh)
        self._detJ0 = self._dm.tangle_init(mesh.coordinates.tensor)
        self._detJ_ratio = Function(mesh.function_space("P0"))

    @property
    def jacobian_determinant_ratio(self):
        """detJ(current) / detJ(initial) per cell, as a P0 field."""
        self._count = self._dm.tangle_check(self.mesh.coordinates.tensor, self._detJ0, self._detJ_ratio.tensor)
        return self._detJ_ratio.written()

    @event("movement.tangling.MeshTanglingChecker.check")
    def check(self):
        """Count cells with ratio <= 0 (owned cells; summed over ranks)."""
        num_tangled = self._dm.tangle_check(self.mesh.coordinates.tensor, self._detJ0, None)
        if self.mesh.part is not None:
            from . import parallel

            num_tangled = int(round(parallel.allreduce_host(num_tangled, "sum")))
        if num_tangled > 0:
            plural = "s" if num_tangled > 1 else ""
            msg = f"Mesh has {num_tangled} tangled element{plural}."
            if self.raise_error:
                raise ValueError(msg)
            warnings.warn(msg, stacklevel=1)
        return num_tangled
