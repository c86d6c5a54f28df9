"""Non-axis-aligned free boundary segments: the EquationBC branch of the L2 projector
(movement/monge_ampere.py:274-309, csrc/mv_eqbc.cu), written to read like the reference's
test/test_monge_ampere.py:113-138 (error cases) and :357-437 (rotated unit domains), with the CPU oracle
(MAOperators.equation_bc_values / l2_project_grad) as the checker."""
import numpy as np
import pytest
import torch

import movement_oracle as mo

pytestmark = pytest.mark.gpu


def _mv():
    import movement_b200 as mv

    mv.mover._print = lambda *a, **k: None
    mv.monge_ampere._print = mv.mover._print
    return mv


def _rotation(dim):
    cs = sn = np.cos(np.pi / 4)
    if dim == 2:
        return np.array([[cs, sn], [-sn, cs]])
    return np.array([[cs, sn, 0.0], [-sn, cs, 0.0], [0.0, 0.0, 1.0]])


def _rotated(dim, n):
    """Unit domain rotated by 45 degrees in the xy-plane (test_monge_ampere.py:388-399): oracle mesh, device mesh."""
    mv = _mv()
    base = mo.unit_square_mesh(n) if dim == 2 else mo.unit_cube_mesh(n)
    R = _rotation(dim)
    om = mo.OMesh(base.coords @ R.T, base.cells, base.bfacets, base.btags)
    return om, mv.Mesh(om.coords, om.cells, om.bfacets, om.btags), R


@pytest.mark.parametrize("dim,n", [(2, 9), (3, 4)])
@pytest.mark.parametrize("fixed", [(), (1,)])
def test_l2_projection_matches_oracle(dim, n, fixed):
    """One projection of a random potential: boundary solve + mass solve against the oracle's sparse-LU answer."""
    mv = _mv()
    from movement_b200 import _lib as L
    import ctypes as C

    om, mesh, _ = _rotated(dim, n)
    mover = mv.MongeAmpereMover(mesh, mv.ring_test_monitor(dim), fixed_boundary_segments=list(fixed))
    ops = mo.MAOperators(om, fixed)
    assert "equation" in ops.bc_kinds
    phi = 0.01 * np.random.default_rng(3).standard_normal(om.num_vertices)
    mover.phi_old.assign(torch.as_tensor(phi))
    mover._sync()
    it = C.c_int(0)
    L.check(L.load().mv_ma_l2_project(mover._ma, C.byref(it)))
    got = mover._grad_phi.tensor.cpu().numpy()
    ref = ops.l2_project_grad(phi)
    assert np.abs(got - ref).max() < 1e-10 * max(1.0, np.abs(ref).max())
    # the boundary values themselves: tangential, zero at the nodes shared by two segments
    gB = ops.equation_bc_values(phi)
    for eq in ops._eq:
        assert np.abs(got[eq["nodes"]] - gB[eq["nodes"]]).max() < 1e-11 * max(1.0, np.abs(gB).max())
        assert np.all(got[eq["corners"]] == 0.0)
        sel = eq["facets"]
        nrm = ops.fnormal[sel][0]
        assert np.abs(got[eq["unknown"]] @ nrm).max() < 1e-12


@pytest.mark.parametrize("dim", [2, 3])
@pytest.mark.parametrize("method", ["relaxation", "quasi_newton"])
@pytest.mark.parametrize("fixed", ["on_boundary", [1], []])
def test_boundary_preservation_non_axis_aligned(dim, method, fixed):
    """test_monge_ampere.py:357-437: the boundaries of the rotated unit domain are preserved, the volume is
    conserved, and the projector carries two EquationBCs per free non-axis-aligned segment."""
    mv = _mv()
    n = 10 if dim == 2 else 5
    om, mesh, R = _rotated(dim, n)
    bnodes = mv.DirichletBC(mesh.coordinates.function_space(), 0, fixed or "on_boundary").nodes
    before = mesh.coordinates.dat.data.copy()
    mover = mv.MongeAmpereMover(mesh, mv.ring_test_monitor(dim), method=method, fixed_boundary_segments=fixed, rtol=1e-3)
    mover.move()
    after = mover.mesh.coordinates.dat.data
    if fixed:
        assert np.allclose(before[bnodes], after[bnodes])
    vol, _ = mo.cell_geometry(after, om.cells)
    assert abs(vol.sum() - 1.0) < 1e-7 and vol.min() > 0
    # rotate back: every boundary vertex is still on its face of the unit domain
    b0, a0 = before @ R, after @ R
    for ax in range(dim):
        for val in (0.0, 1.0):
            on = np.isclose(b0[:, ax], val)
            assert np.allclose(a0[on, ax], val, atol=1e-12)
    assert np.abs(after - before).max() > 1e-3
    bcs = mover._l2_projector._problem.bcs
    n_dir = sum(isinstance(bc, mv.DirichletBC) for bc in bcs)
    n_eq = sum(isinstance(bc, mv.EquationBC) for bc in bcs)
    if fixed == "on_boundary":
        assert (len(bcs), n_dir) == (2 * dim, 2 * dim)
    elif fixed == []:
        assert (len(bcs), n_dir, n_eq) == ((8, 0, 8) if dim == 2 else (10, 2, 8))
    else:
        assert (len(bcs), n_dir, n_eq) == ((7, 1, 6) if dim == 2 else (9, 3, 6))


@pytest.mark.parametrize("dim,n", [(2, 10), (3, 5)])
def test_relaxation_on_rotated_domain_matches_oracle(dim, n):
    """Whole mover on the rotated domain against the oracle: iteration count and coordinates (north_star: +-1, 1e-10)."""
    mv = _mv()
    om, mesh, _ = _rotated(dim, n)
    mover = mv.MongeAmpereMover(mesh, mv.ring_test_monitor(dim), method="relaxation", rtol=1e-3)
    nit = mover.move()
    orc = mo.OracleRelaxation(om, mo.monitor_test_ring, rtol=1e-3)
    ref_nit = orc.move()
    assert nit == ref_nit
    assert np.abs(mover.mesh.coordinates.dat.data - orc.x).max() < 1e-10


def test_non_straight_boundary_valueerror():
    """test_monge_ampere.py:113-120."""
    mv = _mv()
    mesh = mv.UnitSquareMesh(3, 3)
    xy = mesh.coordinates.dat.data
    v = int(np.argmin(np.abs(xy - np.array([0.0, 1.0 / 3.0])).sum(axis=1)))     # a vertex inside segment 1 (x = 0)
    xy[v][0] -= 0.25
    mover = mv.MongeAmpereMover_Relaxation(mesh, mv.ring_test_monitor(2))
    with pytest.raises(ValueError) as cm:
        mover.move()
    assert str(cm.value) == "Boundary segment '1' is not linear."


def test_non_flat_plane_valueerror():
    """test_monge_ampere.py:122-129."""
    mv = _mv()
    mesh = mv.UnitCubeMesh(3, 3, 3)
    xyz = mesh.coordinates.dat.data
    v = int(np.argmin(np.abs(xyz - np.array([0.0, 1.0 / 3.0, 1.0 / 3.0])).sum(axis=1)))   # inside face 1 (x = 0)
    xyz[v][0] -= 0.25
    mover = mv.MongeAmpereMover_Relaxation(mesh, mv.ring_test_monitor(3))
    with pytest.raises(ValueError) as cm:
        mover.move()
    assert str(cm.value) == "Boundary segment '1' is not planar."
