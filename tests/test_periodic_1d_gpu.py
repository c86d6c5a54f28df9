"""SURVEY 8f-2 through the public API: periodic meshes (periodic P1 map + DG coordinate field,
movement/monge_ampere.py:236-239, movement/mover.py:70-75, demos/monge_ampere_periodic.py) and interval meshes
(test/test_monge_ampere.py:174-202,289,322-324) against the CPU oracle, which tests/test_oracle_golden.py pins to the
reference's 90-line periodic trace."""
import json
import os

import numpy as np
import pytest

import movement_oracle as mo

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
with open(os.path.join(HERE, "golden", "demo_traces.json"), encoding="utf-8") as _f:
    GOLD = json.load(_f)["files"]


def _mv():
    import movement_b200 as mv

    return mv


def _lines(capsys):
    return [ln for ln in capsys.readouterr().out.splitlines() if "Volume ratio" in ln]


@pytest.mark.parametrize("dim,n", [(1, 10), (2, 10), (3, 4)])
def test_periodic_relaxation_matches_oracle(dim, n):
    """test/test_monge_ampere.py:296-311: no boundary conditions, domain volume conserved -- plus full parity."""
    mv = _mv()
    mesh = (mv.PeriodicUnitIntervalMesh, mv.PeriodicUnitSquareMesh, mv.PeriodicUnitCubeMesh)[dim - 1](n)
    omesh = (mo.periodic_unit_interval_mesh, mo.periodic_unit_square_mesh, mo.periodic_unit_cube_mesh)[dim - 1](n)
    assert mesh.coordinates.dat.data_ro.shape == omesh.coords.shape and mesh.num_vertices() == omesh.num_vertices
    with pytest.warns(UserWarning, match="no boundary segments"):
        mover = mv.MongeAmpereMover(mesh, mv.ring_test_monitor(dim), method="relaxation", rtol=1e-3)
    assert len(mover._l2_projector._problem.bcs) == 0
    nit = mover.move()
    o = mo.OracleRelaxation(omesh, mo.monitor_test_ring, rtol=1e-3)
    ref_nit = o.move()
    assert nit == ref_nit
    x = mover.mesh.coordinates.dat.data_ro
    assert x.shape == o.x.shape and np.abs(x - o.x).max() < 1e-10
    vol, _ = mo.cell_geometry(x, omesh.ccells)
    assert abs(vol.sum() - 1.0) < 1e-12
    # grad_phi lives in the (DG) coordinate space, _grad_phi in the periodic P1 space (monge_ampere.py:178-179)
    assert mover.grad_phi.dat.data_ro.shape == x.shape
    assert mover._grad_phi.dat.data_ro.shape == (omesh.num_vertices, dim)
    assert np.abs(mover.grad_phi.dat.data_ro - (x - mover.xi.dat.data_ro)).max() < 1e-14
    # the caller's mesh is untouched
    assert np.abs(mesh.coordinates.dat.data_ro - omesh.coords).max() == 0.0


def test_periodic_demo_quasi_newton_trace(capsys):
    """demos/monge_ampere_periodic.py:16-35 through the public API: the first line is the reference's exact known answer,
    the run ends at the reference's fixed point (8.81 / 3.49e-01)."""
    mv = _mv()
    blk = GOLD["demos/monge_ampere_periodic.py"][0]
    mesh = mv.PeriodicUnitSquareMesh(20, 20)
    with pytest.warns(UserWarning, match="no boundary segments"):
        mover = mv.MongeAmpereMover(mesh, mv.ring_test_monitor(2, alpha=20.0, beta=200.0, gamma=0.15), method="quasi_newton",
                                    rtol=1e-8)
    nit = mover.move()
    lines = _lines(capsys)
    assert lines[0].split()[3] == blk["text"][0][0] and lines[0].split()[6] == blk["text"][0][1]
    assert lines[0].split()[-1] == blk["text"][0][2]
    assert abs(nit - blk["converged_in"]) <= 20
    assert lines[-1].split()[3] == blk["text"][-1][0] and lines[-1].split()[6] == blk["text"][-1][1]
    x = mover.mesh.coordinates.dat.data_ro
    vol, _ = mo.cell_geometry(x, mo.periodic_unit_square_mesh(20).ccells)
    assert abs(vol.sum() - 1.0) < 1e-12 and vol.min() > 0.0


@pytest.mark.parametrize("fixed", ["on_boundary", [1], []])
def test_interval_relaxation_matches_oracle(fixed):
    """test/test_monge_ampere.py:322-355 in 1D: every tag of an interval mesh is a fully fixed boundary
    (monge_ampere.py:258-260), whatever fixed_boundary_segments says."""
    mv = _mv()
    n = 10
    mesh = mv.UnitIntervalMesh(n)
    mover = mv.MongeAmpereMover(mesh, mv.ring_test_monitor(1), method="relaxation", rtol=1e-3, fixed_boundary_segments=fixed)
    nit = mover.move()
    o = mo.OracleRelaxation(mo.unit_interval_mesh(n), mo.monitor_test_ring, rtol=1e-3)
    assert nit == o.move()
    x = mover.mesh.coordinates.dat.data_ro
    assert np.abs(x - o.x).max() < 1e-10
    assert x[0, 0] == 0.0 and x[n, 0] == 1.0 and np.abs(x - mesh.coordinates.dat.data_ro).max() > 1e-3
    assert len(mover._l2_projector._problem.bcs) == 2


def test_interval_constant_monitor_and_quasi_newton_error():
    """test/test_monge_ampere.py:174-202 (dim = 1) and :56-62."""
    mv = _mv()
    mesh = mv.UnitIntervalMesh(10)
    mover = mv.MongeAmpereMover(mesh, mv.ConstantMonitorBuilder(1)(), method="relaxation", rtol=1e-3,
                                phi_init=np.zeros(11), H_init=np.zeros((11, 1, 1)))
    assert mover.move() == 0
    assert np.allclose(mesh.coordinates.dat.data_ro, mover.mesh.coordinates.dat.data_ro)
    with pytest.raises(NotImplementedError, match=r"^1D case not implemented for quasi-Newton method\.$"):
        mv.MongeAmpereMover(mesh, mv.ConstantMonitorBuilder(1)(), method="quasi_newton")
