"""Monitor functions -- drop-in for movement/monitor.py:25-171 plus the demo monitors.

A monitor function is a callable ``monitor(mesh)``.  It may return
  * a :class:`BuiltinMonitor` -- evaluated *inside* the coordinate-update kernel (fused; no extra
    pass over the vertices): the Constant/Ball/Ring builders and the three demo monitors;
  * a torch tensor / :class:`~movement_b200.mesh.Function` / numpy array of nodal values [V],
    e.g. computed from ``mesh.coordinates.tensor`` with vectorised torch ops;
  * a python float (constant).
"""
from __future__ import annotations

import abc

from . import _lib as L

__all__ = ["BuiltinMonitor", "MonitorBuilder", "ConstantMonitorBuilder", "BallMonitorBuilder", "RingMonitorBuilder",
           "SolutionBasedMonitorBuilder", "GradientMonitorBuilder", "HessianMonitorBuilder", "GradientHessianMonitorBuilder",
           "helmholtz_hessian_monitor", "sinatan3_hessian_monitor", "ring_test_monitor"]


class BuiltinMonitor:
    """Spec of a monitor evaluated in-kernel: ``kind`` is an ``MV_MON_*`` id, ``params`` its doubles."""

    def __init__(self, kind, params=()):
        self.kind = int(kind)
        self.params = tuple(float(p) for p in params)

    def __eq__(self, other):
        return isinstance(other, BuiltinMonitor) and (self.kind, self.params) == (other.kind, other.params)

    def __hash__(self):
        return hash((self.kind, self.params))

    def __repr__(self):
        return f"BuiltinMonitor(kind={self.kind}, params={self.params})"


class MonitorBuilder(metaclass=abc.ABCMeta):
    """Abstract base class for monitor function factories (movement/monitor.py:25-72)."""

    def __init__(self, dim):
        self.dim = dim

    @abc.abstractmethod
    def _monitor(self, mesh):
        pass  # pragma: no cover

    def get_monitor(self):
        def monitor(mesh):
            return self._monitor(mesh)

        return monitor

    def __call__(self):
        return self.get_monitor()


class ConstantMonitorBuilder(MonitorBuilder):
    """movement/monitor.py:75-89."""

    def _monitor(self, mesh):
        return BuiltinMonitor(L.MV_MON_CONST)


class BallMonitorBuilder(MonitorBuilder):
    r"""m(x) = 1 + alpha / cosh^2(beta ((x-c).(x-c) - gamma^2))  (movement/monitor.py:92-143)."""

    def __init__(self, dim, centre, radius, amplitude, width):
        assert len(centre) == dim
        assert radius > 0
        assert amplitude > 0
        assert width > 0
        super().__init__(dim)
        self.centre = tuple(float(c) for c in centre)
        self.radius = float(radius)
        self.amplitude = float(amplitude)
        self.width = float(width)

    def _monitor(self, mesh):
        c = self.centre + (0.0,) * (3 - len(self.centre))
        return BuiltinMonitor(L.MV_MON_BALL, c + (self.radius, self.amplitude, self.width))


class RingMonitorBuilder(BallMonitorBuilder):
    """movement/monitor.py:146-171."""

    def __init__(self, centre, radius, amplitude, width):
        super().__init__(2, centre, radius, amplitude, width)


class SolutionBasedMonitorBuilder(MonitorBuilder, metaclass=abc.ABCMeta):
    """Abstract base class for monitor factories based on solution data (movement/monitor.py:174-201).

    The recovery operators run on the device (csrc/mv_recovery.cu) on the mesh the monitor is asked for.
    ``projection`` carries the solution onto that mesh: the identity when it has the coordinates of the solution's mesh
    (every test of the reference, test/test_monitor.py), otherwise -- a mover asking for the monitor on its moved copy --
    the Galerkin L2 projection of ``mv_project_p1`` (degree-2 quadrature on the target cells, point location in the
    source mesh).  Firedrake does this through a supermesh; nothing in the reference tree pins that operator."""

    recovery_rtol = 1.0e-12

    @abc.abstractmethod
    def __init__(self, dim, solution):
        from .mesh import Function

        super().__init__(dim)
        assert isinstance(solution, Function)
        self.solution = solution
        self._dm = None
        self._dm_key = None
        self._dm_src = None
        self.projection_outside = 0     # quadrature points of the last projection that fell outside the source mesh

    def projection(self, mesh):
        """movement/monitor.py:191-201."""
        import ctypes as C

        import torch

        from . import parallel
        from .core import device_mesh_from, use_current_stream
        from .mesh import Function

        src = self.solution.space.mesh
        same = mesh is src or (mesh.cells is src.cells or (mesh.cells.shape == src.cells.shape and bool((mesh.cells == src.cells).all())))
        if same and torch.equal(mesh.coordinates.tensor, src.coordinates.tensor):
            out = Function(mesh.function_space(self.solution.space.kind))
            out.assign(self.solution)
            return out
        if self.solution.space.kind != "P1":
            raise NotImplementedError(f"{type(self).__name__}.projection between meshes: only P1 solutions")
        if parallel.is_active() or getattr(mesh, "periodic", False) or getattr(src, "periodic", False):
            raise NotImplementedError(
                f"{type(self).__name__}.projection between meshes with different coordinates is not available on partitioned "
                "or periodic meshes (mv_project_p1 locates points in the whole source mesh on one GPU)")
        if self._dm_src is None:
            self._dm_src = device_mesh_from(src)     # the solution's mesh does not move
        tgt = self._device_mesh(mesh)
        out = Function(mesh.function_space("P1"))
        it, nout = C.c_int(0), C.c_int64(0)
        use_current_stream()
        L.check(L.load().mv_project_p1(self._dm_src.handle, L.ptr(self.solution.tensor), tgt.handle, L.ptr(out.tensor),
                                       self.recovery_rtol, 1000, C.byref(it), C.byref(nout)))
        self.projection_outside = int(nout.value)
        return out.written()

    def _device_mesh(self, mesh):
        """Operators of `mesh` at its current coordinates (ingested again only when the coordinates changed)."""
        import torch

        from .core import device_mesh_from

        x = mesh.coordinates.tensor
        key = (id(mesh), float(x.sum()), float((x * x).sum()))
        if self._dm is None or key != self._dm_key:
            self._dm, self._dm_key = device_mesh_from(mesh), key
        return self._dm

    def _normalised(self, sq):
        """sq / norm(sq, "linf"), the l-infinity norm over every rank (adapt_common.norms.norm)."""
        from . import parallel

        mx = float(sq.abs().max())
        if parallel.is_active():
            mx = parallel.allreduce_host(mx, "max")
        return sq / mx


# TODO of the reference kept: gradient with the Clement interpolant is not offered there either
class GradientMonitorBuilder(SolutionBasedMonitorBuilder):
    r"""m = 1 + alpha grad(u).grad(u) / max grad(u).grad(u)  (movement/monitor.py:204-255)."""

    def __init__(self, dim, solution, scale_factor):
        super().__init__(dim, solution)
        assert scale_factor > 0
        self.gradient_scale_factor = float(scale_factor)

    def recover_gradient(self, target_space):
        """L2-projected gradient of the solution in vector P1 (movement/monitor.py:232-240)."""
        import ctypes as C

        from .core import use_current_stream
        from .mesh import Function

        mesh = target_space.mesh
        u = self.projection(mesh)
        dm = self._device_mesh(mesh)
        g = Function(target_space)
        it = C.c_int(0)
        use_current_stream()
        L.check(L.load().mv_recover_gradient_l2(dm.handle, L.ptr(u.tensor), L.ptr(g.tensor), self.recovery_rtol, 1000, C.byref(it)))
        return g.written()

    def _gradient_part(self, mesh):
        g = self.recover_gradient(mesh.function_space("P1_vec")).tensor
        return self.gradient_scale_factor * self._normalised((g * g).sum(dim=1))

    def _monitor(self, mesh):
        from .mesh import Function

        return Function(mesh.function_space("P1")).assign(1.0 + self._gradient_part(mesh))


# TODO of the reference kept: Hessian by double L2 projection is not offered there either
class HessianMonitorBuilder(SolutionBasedMonitorBuilder):
    r"""m = 1 + beta H(u):H(u) / max H(u):H(u)  (movement/monitor.py:258-308)."""

    def __init__(self, dim, solution, scale_factor):
        super().__init__(dim, solution)
        assert scale_factor > 0
        self.hessian_scale_factor = float(scale_factor)

    def recover_hessian(self, target_space):
        """Hessian by two Clement interpolations (movement/monitor.py:286-293; P1 solutions)."""
        from .core import use_current_stream
        from .mesh import Function

        mesh = target_space.mesh
        u = self.projection(mesh)
        if u.space.kind != "P1":
            raise NotImplementedError("recover_hessian: only P1 solutions (the reference projects P2 and higher first)")
        dm = self._device_mesh(mesh)
        g = Function(mesh.function_space("P1_vec"))
        H = Function(target_space)
        use_current_stream()
        lib = L.load()
        L.check(lib.mv_clement_gradient(dm.handle, 1, L.ptr(u.tensor), L.ptr(g.tensor)))
        L.check(lib.mv_clement_gradient(dm.handle, mesh.geometric_dimension, L.ptr(g.tensor), L.ptr(H.tensor)))
        return H.written()

    def _hessian_part(self, mesh):
        H = self.recover_hessian(mesh.function_space("P1_ten")).tensor
        return self.hessian_scale_factor * self._normalised((H * H).sum(dim=(1, 2)))

    def _monitor(self, mesh):
        from .mesh import Function

        return Function(mesh.function_space("P1")).assign(1.0 + self._hessian_part(mesh))


class GradientHessianMonitorBuilder(GradientMonitorBuilder, HessianMonitorBuilder):
    r"""m = 1 + alpha gg / max gg + beta HH / max HH  (movement/monitor.py:311-365)."""

    def __init__(self, dim, solution, gradient_scale_factor, hessian_scale_factor):
        SolutionBasedMonitorBuilder.__init__(self, dim, solution)
        self.gradient_scale_factor = float(gradient_scale_factor)
        self.hessian_scale_factor = float(hessian_scale_factor)

    def _monitor(self, mesh):
        from .mesh import Function

        return Function(mesh.function_space("P1")).assign(1.0 + self._gradient_part(mesh) + self._hessian_part(mesh))


# ---- demo / test monitors as fused kernels ---------------------------------------------------
def helmholtz_hessian_monitor(alpha=5.0, centre=(0.51, 0.65), width=0.1):
    """demos/monge_ampere_helmholtz.py:54-61,112-122: 1 + alpha |Hess u|_F / max |Hess u|_F,
    u = exp(-|x-c|^2 / w^2) (global max over all ranks)."""
    spec = BuiltinMonitor(L.MV_MON_GAUSS_HESS, (centre[0], centre[1], width, alpha))
    return lambda mesh: spec


def sinatan3_hessian_monitor(alpha=10.0):
    """demos/monge_ampere_3d.py:13-15,49-69: 1 + alpha (H:H)/max(H:H), H = Hess[0.1 sin(50xz) +
    atan2(0.1, sin 5y - 2xz)]."""
    spec = BuiltinMonitor(L.MV_MON_SINATAN3, (alpha,))
    return lambda mesh: spec


def ring_test_monitor(dim, alpha=10.0, beta=200.0, gamma=0.15):
    """test/monitors.py:9-16 (gamma enters unsquared; any dimension)."""
    c = (0.5,) * dim + (0.0,) * (3 - dim)
    spec = BuiltinMonitor(L.MV_MON_BALL_RAW, c + (gamma, alpha, beta))
    return lambda mesh: spec
