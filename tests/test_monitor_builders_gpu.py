"""Solution-based monitor builders (movement/monitor.py:174-365) on the device, written to read like the
reference's test/test_monitor.py:64-152, plus parity of the recovery operators (csrc/mv_recovery.cu) with the
oracle's restatement of animate.recovery on a non-uniform mesh."""
import numpy as np
import pytest
import torch

import movement_oracle as mo

pytestmark = pytest.mark.gpu


def _mv():
    import movement_b200 as mv

    return mv


@pytest.fixture
def setup():
    mv = _mv()
    mesh = mv.UnitSquareMesh(5, 5)
    P1 = mesh.function_space("P1")
    return mv, mesh, mv.Function(P1)


def _x(mesh):
    return mesh.coordinates.tensor[:, 0]


def test_gradient_tiny_scale_factor(setup):
    mv, mesh, solution = setup
    solution.assign(_x(mesh) ** 2)
    mb = mv.GradientMonitorBuilder(dim=2, solution=solution, scale_factor=1e-8)
    assert np.allclose(mb.get_monitor()(mesh).dat.data, 1)


def test_gradient_linear(setup):
    mv, mesh, solution = setup
    solution.assign(_x(mesh))
    mb = mv.GradientMonitorBuilder(dim=2, solution=solution, scale_factor=1)
    assert np.allclose(mb.get_monitor()(mesh).dat.data, 2)


def test_hessian_tiny_scale_factor(setup):
    mv, mesh, solution = setup
    solution.assign(_x(mesh) ** 3)
    mb = mv.HessianMonitorBuilder(dim=2, solution=solution, scale_factor=1e-8)
    assert np.allclose(mb()(mesh).dat.data, 1)


def test_gradient_hessian_consistency(setup):
    """test_monitor.py:102-152: with one tiny factor the combined builder equals the single one."""
    mv, mesh, solution = setup
    solution.assign(_x(mesh) ** 3)
    both = mv.GradientHessianMonitorBuilder(dim=2, solution=solution, gradient_scale_factor=1e-8, hessian_scale_factor=1e-8)
    assert np.allclose(both.get_monitor()(mesh).dat.data, 1)
    mb1 = mv.GradientHessianMonitorBuilder(dim=2, solution=solution, gradient_scale_factor=1, hessian_scale_factor=1e-8)
    mb2 = mv.GradientMonitorBuilder(dim=2, solution=solution, scale_factor=1)
    assert np.abs(mb1.get_monitor()(mesh).dat.data - mb2.get_monitor()(mesh).dat.data).max() < 1e-7
    mb1 = mv.GradientHessianMonitorBuilder(dim=2, solution=solution, gradient_scale_factor=1e-8, hessian_scale_factor=1)
    mb2 = mv.HessianMonitorBuilder(dim=2, solution=solution, scale_factor=1)
    assert np.abs(mb1.get_monitor()(mesh).dat.data - mb2.get_monitor()(mesh).dat.data).max() < 1e-7


@pytest.mark.parametrize("dim,n", [(2, 12), (3, 6), (1, 20)])
def test_recovery_operators_match_oracle(dim, n):
    mv = _mv()
    base = {1: mo.unit_interval_mesh, 2: mo.unit_square_mesh, 3: mo.unit_cube_mesh}[dim](n)
    rng = np.random.default_rng(4)
    c = base.coords.copy()
    interior = np.all((c > 1e-12) & (c < 1 - 1e-12), axis=1)
    c[interior] += 0.3 / n * (rng.random((int(interior.sum()), dim)) - 0.5)
    om = mo.OMesh(c, base.cells, base.bfacets, base.btags)
    mesh = mv.Mesh(om.coords, om.cells, om.bfacets, om.btags)
    u = np.sin(3 * c[:, 0]) + (c ** 2).sum(axis=1)
    solution = mv.Function(mesh.function_space("P1")).assign(torch.as_tensor(u))
    mb = mv.GradientHessianMonitorBuilder(dim=dim, solution=solution, gradient_scale_factor=2.0, hessian_scale_factor=3.0)
    g = mb.recover_gradient(mesh.function_space("P1_vec")).dat.data
    assert np.abs(g - mo.recover_gradient_l2(om, u)).max() < 1e-10
    H = mb.recover_hessian(mesh.function_space("P1_ten")).dat.data
    g_c, H_c = mo.recover_hessian_clement(om, u)
    assert np.abs(H - H_c).max() < 1e-11 * max(1.0, np.abs(H_c).max())
    m = mb.get_monitor()(mesh).dat.data
    assert np.abs(m - mo.solution_monitor(om, u, 2.0, 3.0)).max() < 1e-10


def test_projection_between_different_meshes_is_refused(setup):
    mv, mesh, solution = setup
    solution.assign(_x(mesh))
    other = mv.UnitSquareMesh(5, 5)
    other.coordinates.dat.data[:] *= 1.5
    mb = mv.GradientMonitorBuilder(dim=2, solution=solution, scale_factor=1)
    with pytest.raises(NotImplementedError, match="supermesh"):
        mb.get_monitor()(other)
