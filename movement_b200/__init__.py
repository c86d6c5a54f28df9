"""movement_b200 -- B200-native mesh relocation behind the API of mesh-adaptation/movement.

Drop-in surface (movement/__init__.py:1-6): ``MongeAmpereMover``, ``SpringMover``,
``LaplacianSmoother``, ``MeshTanglingChecker`` and the monitor builders.  The numerical path is
libmovement_b200.so (hand-written sm_100a CUDA behind the C ABI of include/movement_b200.h);
importing the package does not need a GPU, constructing a mover does -- there is no CPU fallback.
"""
from .exceptions import ConvergenceError  # noqa: F401
from .laplacian_smoothing import *  # noqa: F401,F403
from .io import VTKFile  # noqa: F401
from .mesh import (DirichletBC, EquationBC, Function, Mesh, PeriodicUnitCubeMesh, PeriodicUnitIntervalMesh,  # noqa: F401
                   PeriodicUnitSquareMesh, RectangleMesh, UnitCubeMesh, UnitIntervalMesh, UnitSquareMesh,
                   UnitTriangleMesh, from_firedrake)
from .monge_ampere import *  # noqa: F401,F403
from .monitor import *  # noqa: F401,F403
from .spring import *  # noqa: F401,F403
from .tangling import *  # noqa: F401,F403

__version__ = "0.1.0"
