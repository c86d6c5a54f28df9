"""The drop-in Python surface on the GPU, written to read like the reference's own tests
(test/test_monge_ampere.py, test/test_forced_movement.py, test/test_spring.py,
test/test_tangling.py, test/test_demos.py) with the CPU oracle as the checker."""
import json
import math
import os

import numpy as np
import pytest
import torch

import movement_oracle as mo

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
with open(os.path.join(HERE, "golden", "demo_traces.json"), encoding="utf-8") as _f:
    GOLD = json.load(_f)["files"]


def _mv():
    import movement_b200 as mv

    return mv


def _find(coords, pt):
    return int(np.argmin(np.abs(coords - np.asarray(pt, dtype=float)).sum(axis=1)))


def _lines(capsys):
    return [ln for ln in capsys.readouterr().out.splitlines() if "Volume ratio" in ln]


# ------------------------------------------------------------------ errors (test_monge_ampere.py:56-166)
def test_ma_argument_errors():
    mv = _mv()
    mesh = mv.UnitSquareMesh(4, 4)
    with pytest.raises(ValueError, match=r"^Method 'method' not recognised\.$"):
        mv.MongeAmpereMover(mesh, mv.ring_test_monitor(2), method="method")
    with pytest.raises(ValueError, match=r"^Please supply a monitor function\.$"):
        mv.MongeAmpereMover(mesh, None)
    with pytest.raises(NotImplementedError, match=r"^1D case not implemented for quasi-Newton method\.$"):
        mv.MongeAmpereMover(mv.UnitIntervalMesh(4), mv.ring_test_monitor(1), method="quasi_newton")
    with pytest.raises(ValueError, match=r"^Provided boundary_tag '5' is invalid\.$"):
        mv.MongeAmpereMover(mesh, mv.ring_test_monitor(2), fixed_boundary_segments=[5])
    phi_init = mv.Function(mesh.function_space("P1"))
    with pytest.raises(ValueError, match=r"^Need to initialise both phi \*and\* H\.$"):
        mv.MongeAmpereMover_Relaxation(mesh, mv.ring_test_monitor(2), phi_init=phi_init)
    curved = mv.UnitSquareMesh(2, 2)
    curved._degree = 2
    with pytest.raises(NotImplementedError, match=r"^MongeAmpereMover_Relaxation not implemented on curved meshes\.$"):
        mv.MongeAmpereMover(curved, mv.ring_test_monitor(2))


@pytest.mark.parametrize("method", ["relaxation"])
def test_ma_maxiter_and_divergence_messages(method):
    mv = _mv()
    mesh = mv.UnitSquareMesh(4, 4)
    mover = mv.MongeAmpereMover(mesh, mv.ring_test_monitor(2), method=method, maxiter=1)
    with pytest.raises(mv.ConvergenceError, match=r"^Solver failed to converge in 1 iteration\.$"):
        mover.move()
    mover = mv.MongeAmpereMover(mesh, mv.ring_test_monitor(2), method=method, dtol=1.0e-08)
    with pytest.raises(mv.ConvergenceError, match=r"^Solver diverged after 1 iteration\.$"):
        mover.move()
    with pytest.warns(UserWarning):
        mover = mv.MongeAmpereMover(mesh, mv.ring_test_monitor(2), method=method, dtol=1.0e-08,
                                    raise_convergence_errors=False)
    with pytest.raises(Warning, match=r"^Solver diverged after 1 iteration\.$"):
        mover.move()


def test_ma_tangling_valueerror_via_xi():
    """test_monge_ampere.py:148-153: the checker looks at the computational mesh."""
    mv = _mv()
    mover = mv.MongeAmpereMover(mv.UnitSquareMesh(3, 3), mv.ring_test_monitor(2))
    xi = mover.xi.dat.data
    xi[_find(xi, (1 / 3, 1 / 3))] += 0.2
    with pytest.raises(ValueError, match=r"^Mesh has 1 tangled element\.$"):
        mover.move()


# ------------------------------------------------------------------ numerics
@pytest.mark.parametrize("dim", [2, 3])
def test_ma_constant_monitor_zero_iterations(dim):
    """test_monge_ampere.py:174-202."""
    mv = _mv()
    mesh = mv.UnitSquareMesh(5, 5) if dim == 2 else mv.UnitCubeMesh(4, 4, 4)
    coords = mesh.coordinates.dat.data.copy()
    mover = mv.MongeAmpereMover(mesh, mv.ConstantMonitorBuilder(dim)(), rtol=1e-3)
    assert mover.move() == 0
    assert np.allclose(coords, mover.mesh.coordinates.dat.data)
    assert np.allclose(coords, mesh.coordinates.dat.data)  # the caller's mesh is untouched


@pytest.mark.parametrize("dim", [2, 3])
def test_ma_coordinate_update(dim):
    """test_monge_ampere.py:479-503: x = xi + grad(phi) for grad(phi) = e_0."""
    mv = _mv()
    mesh = mv.UnitSquareMesh(2, 2) if dim == 2 else mv.UnitCubeMesh(2, 2, 2)
    mover = mv.MongeAmpereMover(mesh, mv.ConstantMonitorBuilder(dim)())
    mover._grad_phi.interpolate(np.eye(dim)[0])
    mover._update_physical_coordinates()
    mover.to_physical_coordinates()
    xi = mover.xi.dat.data.copy()
    xi[:, 0] += 1
    assert np.abs(mover.x.dat.data - xi).max() == 0.0
    assert np.abs(mover.mesh.coordinates.dat.data - xi).max() == 0.0


def _ref_relax(dim, n, monitor, **kw):
    mesh = mo.unit_square_mesh(n) if dim == 2 else mo.unit_cube_mesh(n)
    o = mo.OracleRelaxation(mesh, monitor, **kw)
    return o, o.move()


@pytest.mark.parametrize("dim,n,expected", [(2, 10, 20), (3, 10, 27)])
def test_ma_relaxation_ring_matches_oracle(dim, n, expected, capsys):
    mv = _mv()
    mesh = mv.UnitSquareMesh(n, n) if dim == 2 else mv.UnitCubeMesh(n, n, n)
    mover = mv.MongeAmpereMover(mesh, mv.ring_test_monitor(dim), method="relaxation", rtol=1e-3)
    nit = mover.move()
    o, ref_nit = _ref_relax(dim, n, mo.monitor_test_ring, rtol=1e-3)
    assert nit == ref_nit == expected                       # iteration count (north_star: +-1)
    assert np.abs(mover.mesh.coordinates.dat.data - o.x).max() < 1e-10   # coordinates (north_star: 1e-10)
    lines = _lines(capsys)
    assert len(lines) == nit + 1
    for ln, (i, vr, cov, res) in zip(lines, o.trace):
        assert ln == f"{i:4d}   Volume ratio {vr:5.2f}   Variation (σ/μ) {cov:8.2e}   Residual {res:8.2e}"


def test_ma_relaxation_helmholtz_demo_full_run(capsys):
    """cfg1-size run to rtol 1e-8: 193 iterations, trace line 0 equals the reference's print
    (demos/monge_ampere_helmholtz.py:158), final state equals its fixed point (:175)."""
    mv = _mv()
    blk = GOLD["demos/monge_ampere_helmholtz.py"][0]
    mover = mv.MongeAmpereMover(mv.UnitSquareMesh(20, 20), mv.helmholtz_hessian_monitor(), method="relaxation", rtol=1e-8)
    nit = mover.move()
    o, ref_nit = _ref_relax(2, 20, mo.monitor_helmholtz_exact, rtol=1e-8)
    assert abs(nit - ref_nit) <= 1 and ref_nit == 193
    assert np.abs(mover.mesh.coordinates.dat.data - o.x).max() < 1e-10
    lines = _lines(capsys)
    assert lines[0].split()[3] == blk["text"][0][0] and lines[0].split()[-1] == blk["text"][0][2]
    assert lines[-1].split()[3] == blk["text"][-1][0] and lines[-1].split()[6] == blk["text"][-1][1]


def test_ma_torch_callable_monitor_equals_fused_monitor():
    """A vectorised torch monitor (evaluated by the caller on mesh.coordinates.tensor) and the fused
    in-kernel monitor give the same mesh."""
    mv = _mv()

    def ring(mesh):
        x = mesh.coordinates.tensor
        r = ((x - 0.5) ** 2).sum(dim=1)
        return 1.0 + 10.0 / torch.cosh(200.0 * (r - 0.15)) ** 2

    a = mv.MongeAmpereMover(mv.UnitSquareMesh(10, 10), ring, rtol=1e-3)
    b = mv.MongeAmpereMover(mv.UnitSquareMesh(10, 10), mv.ring_test_monitor(2), rtol=1e-3)
    assert a.move() == b.move() == 20
    assert np.abs(a.mesh.coordinates.dat.data - b.mesh.coordinates.dat.data).max() < 1e-11


def test_ma_continue_and_change_monitor():
    """test_monge_ampere.py:204-254: a second move() continues from phi/H; switching to a constant
    monitor returns towards the original mesh."""
    mv = _mv()
    mesh = mv.UnitSquareMesh(10, 10)
    coords = mesh.coordinates.dat.data.copy()
    mover = mv.MongeAmpereMover(mesh, mv.ring_test_monitor(2), rtol=1e-1)
    n1 = mover.move()
    mover.rtol = 1e-3
    n2 = mover.move()
    full = mv.MongeAmpereMover(mesh, mv.ring_test_monitor(2), rtol=1e-3)
    n_full = full.move()
    assert n2 <= n_full and n1 + n2 >= n_full - 1
    assert np.abs(mover.mesh.coordinates.dat.data - full.mesh.coordinates.dat.data).max() < 2e-3
    assert not np.allclose(coords, mover.mesh.coordinates.dat.data)
    mover.monitor_function = mv.ConstantMonitorBuilder(2)()
    mover.move()
    assert np.abs(mover.mesh.coordinates.dat.data - coords).max() < 1e-3


@pytest.mark.parametrize("dim", [2, 3])
@pytest.mark.parametrize("fixed", ["on_boundary", [1], []])
def test_ma_boundary_preservation_axis_aligned(dim, fixed):
    """test_monge_ampere.py:318-355."""
    mv = _mv()
    n = 10 if dim == 2 else 5
    mesh = mv.UnitSquareMesh(n, n) if dim == 2 else mv.UnitCubeMesh(n, n, n)
    bnodes = mv.DirichletBC(mesh.coordinates.function_space(), 0, fixed or "on_boundary").nodes
    before = mesh.coordinates.dat.data.copy()
    mover = mv.MongeAmpereMover(mesh, mv.ring_test_monitor(dim), fixed_boundary_segments=fixed, rtol=1e-3)
    mover.move()
    after = mover.mesh.coordinates.dat.data
    if fixed:
        assert np.allclose(before[bnodes], after[bnodes])
    vol, _ = mo.cell_geometry(after, mesh.cells)
    assert abs(vol.sum() - 1.0) < 1e-7
    allb = mv.DirichletBC(mesh.coordinates.function_space(), 0, "on_boundary").nodes
    for ax in range(dim):   # every boundary vertex stays on its bounding plane
        on0, on1 = before[allb, ax] == 0.0, before[allb, ax] == 1.0
        assert np.allclose(after[allb][on0, ax], 0.0, atol=1e-13) and np.allclose(after[allb][on1, ax], 1.0, atol=1e-13)
    bcs = mover._l2_projector._problem.bcs
    assert len(bcs) == 2 * dim and all(isinstance(bc, mv.DirichletBC) for bc in bcs)


# ------------------------------------------------------------------ tangling (test_tangling.py)
def test_tangling_checker_api():
    mv = _mv()
    checker = mv.MeshTanglingChecker(mv.UnitSquareMesh(3, 3), raise_error=False)
    assert checker.check() == 0
    c = checker.mesh.coordinates.dat.data
    c[_find(c, (1 / 3, 1 / 3))] += 0.2
    with pytest.warns(UserWarning, match="Mesh has 1 tangled element."):
        assert checker.check() == 1
    ratio = checker.jacobian_determinant_ratio.dat.data
    assert (ratio <= 0).sum() == 1
    checker = mv.MeshTanglingChecker(mv.UnitSquareMesh(3, 3), raise_error=True)
    c = checker.mesh.coordinates.dat.data
    c[_find(c, (1 / 3, 1 / 3))] += 0.5
    with pytest.raises(ValueError, match=r"^Mesh has 3 tangled elements\.$"):
        checker.check()
    checker = mv.MeshTanglingChecker(mv.UnitCubeMesh(3, 3, 3), raise_error=False)
    c = checker.mesh.coordinates.dat.data
    c[_find(c, (1 / 3, 1 / 3, 1 / 3)), 0] += 0.4
    with pytest.warns(UserWarning):
        assert checker.check() == 6


def test_tangling_checker_1d():
    """test/test_tangling.py:14-37 (interval meshes carry geometry only; the movers reject them)."""
    mv = _mv()
    checker = mv.MeshTanglingChecker(mv.UnitIntervalMesh(3), raise_error=True)
    checker.mesh.coordinates.dat.data[1] += 0.4
    with pytest.raises(ValueError, match=r"^Mesh has 1 tangled element\.$"):
        checker.check()
    checker = mv.MeshTanglingChecker(mv.UnitIntervalMesh(3), raise_error=False)
    assert checker.check() == 0
    checker.mesh.coordinates.dat.data[1] += 0.4
    with pytest.warns(UserWarning, match="Mesh has 1 tangled element."):
        assert checker.check() == 1
    checker = mv.MeshTanglingChecker(mv.UnitIntervalMesh(3), raise_error=False)
    checker.mesh.coordinates.dat.data[1] += 0.4
    checker.mesh.coordinates.dat.data[3] -= 0.4
    with pytest.warns(UserWarning, match="Mesh has 2 tangled elements."):
        assert checker.check() == 2
    checker = mv.MeshTanglingChecker(mv.UnitIntervalMesh(3), raise_error=False)
    checker.mesh.coordinates.interpolate(lambda x: 1.0 - x)          # flip: every element inverts
    with pytest.warns(UserWarning):
        assert checker.check() == checker.mesh.num_cells()
    ratio = checker.jacobian_determinant_ratio.dat.data
    assert np.allclose(ratio, -1.0, rtol=1e-14, atol=0.0)


# ------------------------------------------------------------------ forced movement (test_forced_movement.py)
def _move(mv, cls, mesh, fixed=None, moving=None, vector=None, **kwargs):
    mover = cls(mesh, 1.0)
    bcs = []
    if fixed:
        bcs.append(mv.DirichletBC(mover.coord_space, 0, fixed))
    if moving:
        f = mv.Function(mover.coord_space).interpolate(vector or [1, 0])
        bcs.append(mv.DirichletBC(mover.coord_space, f, moving))
    mover.move(0.0, boundary_conditions=bcs, **kwargs)
    return mover.mesh


def _movers():
    mv = _mv()
    return [mv.LaplacianSmoother, mv.SpringMover]


@pytest.mark.parametrize("which", [0, 1])
def test_forced_movement_properties(which):
    mv = _mv()
    cls = _movers()[which]
    for mesh in (mv.UnitTriangleMesh(), mv.UnitSquareMesh(1, 1)):
        coords = mesh.coordinates.dat.data.copy()
        assert np.allclose(coords, _move(mv, cls, mesh).coordinates.dat.data)
    mesh = mv.UnitSquareMesh(10, 10)
    coords = mesh.coordinates.dat.data.copy()
    new = _move(mv, cls, mesh, moving=1)
    assert np.allclose(new.coordinates.dat.data, coords + [1.0, 0.0])      # rigid shift by exactly 1
    back = _move(mv, cls, new, moving=1, vector=[-1, 0])
    assert np.allclose(back.coordinates.dat.data, coords)                  # forward + back
    stretched = _move(mv, cls, mesh, moving=1, fixed=2, vector=[0.1, 0]).coordinates.dat.data
    assert not np.allclose(stretched, coords)


@pytest.mark.parametrize("which", [0, 1])
def test_forced_movement_time_dependent_bc(which):
    """test_forced_movement.py:108-123,146-161."""
    mv = _mv()
    mesh = mv.UnitSquareMesh(10, 10)
    coords = mesh.coordinates.dat.data.copy()
    mover = _movers()[which](mesh, 1.0)
    forcing = mv.Function(mover.coord_space)
    bc = mv.DirichletBC(mover.coord_space, forcing, 1)

    def update_bc(time):
        forcing.interpolate([math.cos(math.pi * time / 2), math.sin(math.pi * time / 2)])

    kw = "update_boundary_velocity" if which == 0 else "update_boundary_displacement"
    mover.move(0.0, boundary_conditions=bc, **{kw: update_bc})
    assert not np.allclose(coords, mover.mesh.coordinates.dat.data)
    assert np.allclose(coords, mesh.coordinates.dat.data)


def test_laplacian_convergence_error():
    """test_forced_movement.py:135-144."""
    mv = _mv()
    with pytest.raises(mv.ConvergenceError, match=r"^Solver failed to converge\.$"):
        _move(mv, mv.LaplacianSmoother, mv.UnitSquareMesh(10, 10), moving=1, fixed=2, vector=[1.0, 0],
              solver_parameters={"ksp_type": "cg", "ksp_max_it": 0})


def test_laplacian_tangling_error():
    """test_laplacian_smoothing.py:9-20 (a forcing large enough to invert cells)."""
    mv = _mv()
    mesh = mv.UnitSquareMesh(3, 3)
    mover = mv.LaplacianSmoother(mesh, 1.0)
    f = mv.Function(mover.coord_space).interpolate([0.0, -2.0])
    bcs = [mv.DirichletBC(mover.coord_space, 0, [1, 2, 3]), mv.DirichletBC(mover.coord_space, f, 4)]
    with pytest.raises(ValueError, match=r"^Mesh has \d+ tangled elements?\.$"):
        mover.move(0.0, boundary_conditions=bcs)


def _demo_forced(mv, cls, capsys, read_moved):
    n = 10
    mesh = mv.UnitSquareMesh(n, n)
    mover = cls(mesh, 0.1)
    top = mv.Function(mover.coord_space)
    moving = mv.DirichletBC(mover.coord_space, top, 4)

    def update(t):
        src = mover.mesh if read_moved else mesh
        coord = src.coordinates.dat.data_ro
        data = top.dat.data
        for i in moving.nodes:
            data[i][1] = 0.2 * math.sin(2 * math.pi * t) * math.sin(math.pi * coord[i][0])

    fixed = mv.DirichletBC(mover.coord_space, 0, [1, 2, 3])
    kw = "update_boundary_displacement" if read_moved else "update_boundary_velocity"
    for k in range(11):
        mover.move(0.1 * k, boundary_conditions=(fixed, moving), **{kw: update})
    return mesh, mover, _lines(capsys)


def test_laplacian_demo_trace(capsys):
    """demos/laplacian_smoothing.py:161-171 through the public API."""
    mv = _mv()
    blk = GOLD["demos/laplacian_smoothing.py"][0]
    mesh, mover, lines = _demo_forced(mv, mv.LaplacianSmoother, capsys, read_moved=False)
    assert len(lines) == 11
    for k, ln in enumerate(lines):
        tok = ln.split()
        assert tok[0] == f"{0.1 * k:.2f}" and tok[1] == "s"
        assert tok[4] == blk["text"][k][0] and tok[-2] == blk["text"][k][2]
        if k not in (0, 9, 10):
            assert tok[7] == blk["text"][k][1]
    assert np.abs(mover.mesh.coordinates.dat.data - mesh.coordinates.dat.data).max() < 1e-10


def test_spring_demo_trace(capsys):
    """demos/lineal_spring.py:176-186,204 through the public API."""
    mv = _mv()
    blk = GOLD["demos/lineal_spring.py"][0]
    mesh, mover, lines = _demo_forced(mv, mv.SpringMover, capsys, read_moved=True)
    assert len(lines) == 11
    for k, ln in enumerate(lines):
        tok = ln.split()
        assert tok[3] == blk["text"][k][0] and tok[-2] == blk["text"][k][2]
        if k > 0:
            assert tok[6] == blk["text"][k][1]
    linf = np.abs(mover.mesh.coordinates.dat.data - mesh.coordinates.dat.data).max()
    assert f"{linf:.3f}" == "0.001"
    K = mover.assemble_stiffness_matrix(boundary_conditions=None)
    assert K.shape == (242, 242)


# ------------------------------------------------------------------ spring internals (test_spring.py)
def test_spring_errors_and_quantities():
    mv = _mv()
    with pytest.raises(ValueError, match=r"^Method 'method' not recognised\.$"):
        mv.SpringMover(mv.UnitSquareMesh(1, 1), 1.0, method="method")
    with pytest.raises(NotImplementedError, match=r"^Torsional springs not yet implemented\.$"):
        mv.SpringMover(mv.UnitSquareMesh(1, 1), 1.0, method="torsional")
    with pytest.raises(ValueError, match=r"^Spring-based Movers only currently support triangular meshes\.$"):
        mv.SpringMover(mv.UnitCubeMesh(1, 1, 1), 1.0)
    mover = mv.SpringMover(mv.UnitTriangleMesh(), 1.0)
    with pytest.raises(ValueError, match=r"contains invalid boundary tags\.$"):
        mover.assemble_stiffness_matrix(mv.DirichletBC(mover.coord_space, 0, 2))
    other = mv.UnitTriangleMesh()
    with pytest.raises(ValueError, match=r"coord_space as their function space\.$"):
        mover.assemble_stiffness_matrix(mv.DirichletBC(other.coordinates.function_space(), 0, 1))
    # test_spring.py:136-157 -- unit right triangle
    edges = [tuple(e) for e in mover.edges.tolist()]
    L_, T_, A_ = mover.facet_areas.dat.data, mover.tangents.dat.data, np.degrees(mover.angles.dat.data)
    k = edges.index((0, 1)); assert np.isclose(L_[k], 1) and np.allclose(np.abs(T_[k]), [1, 0]) and np.isclose(A_[k] % 180, 0)
    k = edges.index((1, 2)); assert np.isclose(L_[k], math.sqrt(2)) and np.isclose(A_[k], 135)
    k = edges.index((0, 2)); assert np.isclose(L_[k], 1) and np.allclose(np.abs(T_[k]), [0, 1]) and np.isclose(A_[k], 90)


@pytest.mark.parametrize("which", ["triangle", "square"])
def test_spring_stiffness_boundary_conditions(which):
    """test_spring.py:77-120."""
    mv = _mv()
    mesh = mv.UnitTriangleMesh() if which == "triangle" else mv.UnitSquareMesh(1, 1)
    mover = mv.SpringMover(mesh, 1.0, method="lineal")
    ident = np.eye(2 * mesh.num_vertices())
    K = mover._stiffness_matrix()
    assert not np.allclose(K, ident)
    K_bc = mover.assemble_stiffness_matrix()
    assert not np.allclose(K, K_bc) and np.allclose(K_bc, ident)
    # the device operator is the same matrix
    import ctypes as C

    from movement_b200 import _lib as L

    mesh2 = mv.UnitSquareMesh(4, 4)
    mover = mv.SpringMover(mesh2, 1.0)
    bc = mv.DirichletBC(mover.coord_space, 0, 1)
    Kd = mover.assemble_stiffness_matrix(bc)
    x = np.random.default_rng(0).standard_normal(2 * mesh2.num_vertices())
    y = np.empty_like(x)
    mask = mover._mask.copy()
    L.check(L.load().mv_spring_apply(mover._dm.handle, L.ptr(mover.mesh.coordinates.tensor), L.ptr(mask), L.ptr(x), L.ptr(y)))
    assert np.abs(y - Kd @ x).max() < 1e-12


# ------------------------------------------------------------------ quasi-Newton
def test_qn_residual_and_jacobian_match_oracle():
    """F and J from the device against the numpy oracle's func()/jacobian() at a non-trivial state."""
    import ctypes as C

    from movement_b200 import _lib as L

    mv = _mv()
    n = 8
    mover = mv.MongeAmpereMover(mv.UnitSquareMesh(n, n), mv.helmholtz_hessian_monitor(), method="quasi_newton")
    omesh = mo.unit_square_mesh(n)
    q = mo.OracleQuasiNewton(omesh, mo.monitor_helmholtz_exact)
    V = omesh.num_vertices
    rng = np.random.default_rng(3)
    phi = 1e-3 * rng.standard_normal(V)
    H = 5e-2 * rng.standard_normal((V, 2, 2))
    mover.phi.assign(phi); mover.H.assign(H)
    mover._begin()
    f2 = mover._func()
    u = np.concatenate([phi, H.ravel()])
    Fref = q.func(u)
    F = mover._F.cpu().numpy()
    assert abs(f2 - Fref @ Fref) < 1e-10 * (Fref @ Fref)
    assert np.abs(F - Fref).max() < 1e-12 * max(1.0, np.abs(Fref).max())
    J = q.jacobian(u)
    z = rng.standard_normal(u.size)
    z[:V] -= z[:V].mean()
    y = np.empty_like(z)
    zt = torch.as_tensor(z).cuda(); yt = torch.empty_like(zt)
    L.check(L.load().mv_qn_jacobian_apply(mover._ma, L.ptr(zt), L.ptr(yt)))
    ref = J @ z
    ref[:V] -= ref[:V].mean()
    assert np.abs(yt.cpu().numpy() - ref).max() < 1e-11 * np.abs(ref).max()


@pytest.mark.parametrize("dim,n", [(2, 20), (3, 6)])
def test_qn_converges_to_the_oracle_fixed_point(dim, n, capsys):
    """Both Newton iterations are driven to ||F|| < 2e-16 (the map F -> x amplifies by ~1e3 at this h); the moved meshes must then coincide
    (north_star: coordinates within 1e-10), and the 2D run reproduces the reference's 17 SNES
    iterations at rtol 1e-8 (demos/monge_ampere_helmholtz.py:175-177)."""
    mv = _mv()
    mesh = mv.UnitSquareMesh(n, n) if dim == 2 else mv.UnitCubeMesh(n, n, n)
    mon = mv.helmholtz_hessian_monitor() if dim == 2 else mv.ring_test_monitor(3)
    ref_mon = mo.monitor_helmholtz_exact if dim == 2 else mo.monitor_test_ring
    mover = mv.MongeAmpereMover(mesh, mon, method="quasi_newton", rtol=1e-8)
    mover.ksp_rtol = 1e-9
    nit = mover.move()
    omesh = mo.unit_square_mesh(n) if dim == 2 else mo.unit_cube_mesh(n)
    q = mo.OracleQuasiNewton(omesh, ref_mon, rtol=1e-8)
    ref_nit = q.move()
    assert abs(nit - ref_nit) <= 1
    if dim == 2:
        assert ref_nit == 17
        lines = _lines(capsys)
        blk = GOLD["demos/monge_ampere_helmholtz.py"][0]
        assert lines[0].split()[3] == blk["text"][0][0] and lines[-1].split()[6] == blk["text"][-1][1]
    # tighten both to the fixed point
    mover.rtol = 2e-16; mover.snes_rtol = 0.0; mover.maxiter = 40
    mover.move()
    q2 = mo.OracleQuasiNewton(omesh, ref_mon, rtol=2e-16, maxiter=60, snes_rtol=0.0)
    q2.move()
    err = np.abs(mover.mesh.coordinates.dat.data - q2.x).max()
    assert err < 1e-10, err


def test_qn_maxiter_message():
    mv = _mv()
    mover = mv.MongeAmpereMover(mv.UnitSquareMesh(4, 4), mv.ring_test_monitor(2), method="quasi_newton", maxiter=1)
    with pytest.raises(mv.ConvergenceError, match=r"^Solver failed to converge in 1 iteration\.$"):
        mover.move()


def test_qn_linear_solve_failure_is_a_convergence_error():
    """ksp_max_it = 1: GMRES cannot reach ksp_rtol -> KSP_DIVERGED_ITS -> SNES_DIVERGED_LINEAR_SOLVE, which the reference
    turns into ConvergenceError with the SNES iteration count (movement/monge_ampere.py:744-745); here zero Newton
    iterations were completed, so the message has no count."""
    mv = _mv()
    mover = mv.MongeAmpereMover(mv.UnitSquareMesh(8, 8), mv.ring_test_monitor(2), method="quasi_newton")
    mover.ksp_max_it = 1
    mover.ksp_rtol = 1e-12
    with pytest.raises(mv.ConvergenceError, match=r"^Solver failed to converge\.$"):
        mover.move()
    assert mover.ksp_iterations == [1] and mover.snes_iterations == 0
    # the reference's restart length (solver_parameters.py:13) is the default and fits at this size
    mover2 = mv.MongeAmpereMover(mv.UnitSquareMesh(8, 8), mv.ring_test_monitor(2), method="quasi_newton", rtol=1e-3)
    assert mover2.ksp_gmres_restart == 200
    mover2.move()
    assert mover2.gmres_restart_used == 200


def test_spring_dirichlet_rows_follow_the_current_call():
    """spring.py:183-235 rebuilds K on every call: only the boundary conditions of THIS call replace rows, while the
    forcing written by earlier calls persists (spring.py:229-230).  Two consecutive moves with different tag sets
    against the oracle's dense row-replaced system."""
    mv = _mv()
    n = 6
    mesh = mv.UnitSquareMesh(n, n)
    mover = mv.SpringMover(mesh, 1.0)
    fA = mv.Function(mover.coord_space).interpolate([0.0, 0.01])
    bcA = [mv.DirichletBC(mover.coord_space, fA, 4), mv.DirichletBC(mover.coord_space, 0, [1, 2, 3])]
    bcB = [mv.DirichletBC(mover.coord_space, 0, [1, 2])]
    omesh = mo.unit_square_mesh(n)
    ref = mo.OracleSpringMover(omesh, 1.0)
    nodesA4 = np.unique(omesh.bfacets[omesh.btags == 4])
    nodesA0 = np.unique(omesh.bfacets[np.isin(omesh.btags, [1, 2, 3])])
    # call 1: sets A (later BCs override earlier ones at shared vertices)
    mover.move(0.0, boundary_conditions=bcA)
    ref.move(np.concatenate([nodesA4, nodesA0]),
             np.concatenate([np.tile([0.0, 0.01], (len(nodesA4), 1)), np.zeros((len(nodesA0), 2))]))
    assert np.abs(mover.mesh.coordinates.dat.data_ro - ref.x).max() < 1e-10
    # call 2: set B pins only x = 0 and x = 1; the top vertices keep their forcing but are free rows now
    nodesB = np.unique(omesh.bfacets[np.isin(omesh.btags, [1, 2])])
    mover.move(1.0, boundary_conditions=bcB)
    ref.move(nodesB, np.zeros((len(nodesB), 2)))
    assert np.abs(mover.mesh.coordinates.dat.data_ro - ref.x).max() < 1e-10
    free_top = np.setdiff1d(nodesA4, nodesB)
    assert len(free_top) and np.abs(mover.displacement.reshape(-1, 2)[free_top]).max() > 1e-6
