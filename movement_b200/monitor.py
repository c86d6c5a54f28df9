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
           "GradientMonitorBuilder", "HessianMonitorBuilder", "GradientHessianMonitorBuilder",
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


class _SolutionBased(MonitorBuilder):
    """movement/monitor.py:174-365 need ``animate``'s recovery operators and cross-mesh projection
    (marked 'next'); not on the accelerated path yet."""

    def __init__(self, *args, **kwargs):
        raise NotImplementedError(
            f"{type(self).__name__}: solution-based monitors are outside the accelerated path "
            "(SURVEY 8f-3); pass a torch callable that returns nodal monitor values instead.")

    def _monitor(self, mesh):  # pragma: no cover
        raise NotImplementedError


class GradientMonitorBuilder(_SolutionBased):
    pass


class HessianMonitorBuilder(_SolutionBased):
    pass


class GradientHessianMonitorBuilder(_SolutionBased):
    pass


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
