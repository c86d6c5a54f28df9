"""Parity of every CUDA kernel group against the CPU oracle, through the C ABI."""
import numpy as np
import pytest
import torch

import movement_oracle as mo

pytestmark = pytest.mark.gpu


def _mesh(dim, n):
    return mo.unit_square_mesh(n) if dim == 2 else mo.unit_cube_mesh(n)


def _perturbed(mesh, amp=0.2):
    """Interior vertices jiggled so that no two cells are congruent."""
    rng = np.random.default_rng(1)
    c = mesh.coords.copy()
    h = 1.0 / round(mesh.num_vertices ** (1.0 / mesh.dim) - 1)
    interior = np.all((c > 1e-12) & (c < 1 - 1e-12), axis=1)
    c[interior] += amp * h * (rng.random((interior.sum(), mesh.dim)) - 0.5)
    return mo.OMesh(c, mesh.cells, mesh.bfacets, mesh.btags)


def _dmesh(mesh):
    from movement_b200.core import DeviceMesh

    return DeviceMesh(mesh.coords, mesh.cells, mesh.bfacets, mesh.btags)


CASES = [(2, 7), (2, 20), (3, 3), (3, 6)]


@pytest.mark.parametrize("dim,n", CASES)
def test_ingest_operators(dim, n):
    from movement_b200 import _lib as L

    mesh = _perturbed(_mesh(dim, n))
    dm = _dmesh(mesh)
    ops = mo.MAOperators(mesh)
    M = dm.to_scipy(L.MV_GET_MASS)
    K = dm.to_scipy(L.MV_GET_STIFF)
    assert abs(M - ops.M).max() < 1e-15
    assert abs(K - ops.K).max() < 1e-12 * abs(ops.K).max()
    assert (M != 0).sum() == (ops.M != 0).sum() == dm.nnz
    np.testing.assert_allclose(dm.get(L.MV_GET_LUMPED), ops.ml, rtol=1e-13)
    np.testing.assert_allclose(dm.get(L.MV_GET_CELL_VOL), ops.vol, rtol=1e-13)
    np.testing.assert_allclose(dm.total_volume, ops.total_volume, rtol=1e-13)
    np.testing.assert_allclose(dm.ml_sumsq, (ops.ml ** 2).sum(), rtol=1e-13)
    # boundary facets
    np.testing.assert_array_equal(dm.get(L.MV_GET_BF_CELL), ops.fcell)
    np.testing.assert_allclose(dm.get(L.MV_GET_BF_NORMAL), ops.fnormal, atol=1e-14)
    np.testing.assert_allclose(dm.get(L.MV_GET_BF_AREA), ops.farea, rtol=1e-13)


@pytest.mark.parametrize("dim,n", [(2, 9), (3, 4)])
@pytest.mark.parametrize("ncomp", [1, 2, 3, 4, 9])
def test_spmv(dim, n, ncomp):
    from movement_b200 import _lib as L

    if ncomp > dim * dim and ncomp != 9:
        pytest.skip("n/a")
    mesh = _perturbed(_mesh(dim, n))
    dm = _dmesh(mesh)
    ops = mo.MAOperators(mesh)
    rng = np.random.default_rng(2)
    x = rng.standard_normal((mesh.num_vertices, ncomp))
    for op, A in ((L.MV_OP_MASS, ops.M), (L.MV_OP_STIFFNESS, ops.K), (L.MV_OP_MASS_PLUS_STIFFNESS, ops.M + ops.K)):
        xd = torch.as_tensor(x).cuda()
        if ncomp == 1:
            xd = xd[:, 0].contiguous()
        y = dm.spmv(op, xd).cpu().numpy().reshape(mesh.num_vertices, ncomp)
        ref = A @ x
        assert np.abs(y - ref).max() < 1e-13 * max(1.0, np.abs(ref).max())


@pytest.mark.parametrize("dim,n", [(2, 12), (3, 5)])
def test_pcg_mass_multi_rhs_and_mask(dim, n):
    from movement_b200 import _lib as L

    mesh = _perturbed(_mesh(dim, n))
    dm = _dmesh(mesh)
    ops = mo.MAOperators(mesh)
    V = mesh.num_vertices
    rng = np.random.default_rng(3)
    # d*d components, no mask (A8 solve)
    b = rng.standard_normal((V, dim * dim))
    x = dm.zeros(V, dim * dim)
    it, rel = dm.pcg(L.MV_OP_MASS, torch.as_tensor(b).cuda(), x, rtol=1e-13)
    ref = ops.solve_mass(b)
    assert it < 200 and rel.max() <= 1e-13
    assert np.abs(x.cpu().numpy() - ref).max() < 1e-10 * np.abs(ref).max()
    # d components with the axis-aligned Dirichlet mask (A1 solve)
    b = rng.standard_normal((V, dim))
    mask = (ops.bc_mask * (1 << np.arange(dim))).sum(axis=1).astype(np.uint16)
    x = dm.zeros(V, dim)
    it, rel = dm.pcg(L.MV_OP_MASS, torch.as_tensor(b).cuda(), x, mask=torch.as_tensor(mask.view(np.int16)).cuda(), rtol=1e-13)
    ref = np.stack([ops.solve_mass_bc(b[:, a], a) for a in range(dim)], axis=1)
    assert np.abs(x.cpu().numpy() - ref).max() < 1e-10 * np.abs(ref).max()
    assert np.all(x.cpu().numpy()[ops.bc_mask] == 0.0)


@pytest.mark.parametrize("dim,n", [(2, 16), (3, 6)])
def test_pcg_poisson_nullspace(dim, n):
    from movement_b200 import _lib as L

    mesh = _perturbed(_mesh(dim, n))
    dm = _dmesh(mesh)
    ops = mo.MAOperators(mesh)
    V = mesh.num_vertices
    rng = np.random.default_rng(4)
    b = rng.standard_normal(V) * ops.ml
    x = dm.zeros(V)
    it, rel = dm.pcg(L.MV_OP_STIFFNESS, torch.as_tensor(b).cuda(), x, rtol=1e-13, project_mean=True)
    ref = ops.solve_poisson_nullspace(b)
    ref -= ref.mean()
    got = x.cpu().numpy()
    assert abs(got.mean()) < 1e-14
    assert np.abs(got - ref).max() < 1e-9 * np.abs(ref).max()


def test_pcg_maxit_reports_not_converged():
    from movement_b200 import _lib as L

    mesh = _mesh(2, 16)
    dm = _dmesh(mesh)
    b = torch.ones(mesh.num_vertices, dtype=torch.float64, device="cuda")
    b -= b.mean() + torch.linspace(-1, 1, mesh.num_vertices, dtype=torch.float64, device="cuda")
    x = dm.zeros(mesh.num_vertices)
    dm.set_preconditioner("jacobi")   # (the 289-vertex mesh is solved directly by the multigrid coarse solve)
    with pytest.raises(L.MovementB200Error) as e:
        dm.pcg(L.MV_OP_STIFFNESS, b, x, maxit=3, project_mean=True)
    assert e.value.status == L.MV_ERR_NOT_CONVERGED


@pytest.mark.parametrize("dim,n", [(2, 9), (3, 4)])
def test_ma_assembly_kernels(dim, n):
    import ctypes as C

    from movement_b200 import _lib as L

    mesh = _perturbed(_mesh(dim, n))
    dm = _dmesh(mesh)
    ops = mo.MAOperators(mesh)
    V = mesh.num_vertices
    rng = np.random.default_rng(5)
    phi = rng.standard_normal(V)
    lib = L.load()
    b = np.empty((V, dim))
    L.check(lib.mv_ma_l2_rhs(dm.handle, L.ptr(phi), L.ptr(b)))
    ref = ops.l2_rhs(phi)
    assert np.abs(b - ref).max() < 1e-13 * np.abs(ref).max()
    c = np.empty((V, dim, dim))
    L.check(lib.mv_ma_hess_rhs(dm.handle, L.ptr(phi), L.ptr(c)))
    ref = ops.hess_rhs(phi)
    assert np.abs(c - ref).max() < 1e-13 * np.abs(ref).max()
    m = 1.0 + rng.random(V)
    H = 0.3 * rng.standard_normal((V, dim, dim))
    a = np.empty(V)
    L.check(lib.mv_ma_assemble_a(dm.handle, L.ptr(m), L.ptr(H), L.ptr(a)))
    ref = ops.assemble_a(m, H)
    assert np.abs(a - ref).max() < 1e-13 * np.abs(ref).max()


@pytest.mark.parametrize("dim,n", [(2, 9), (3, 4)])
def test_bc_mask_axis_aligned(dim, n):
    from movement_b200.core import MAState

    mesh = _mesh(dim, n)
    dm = _dmesh(mesh)
    st = MAState(dm, mesh.coords)
    st.set_boundary(())
    ops = mo.MAOperators(mesh)
    ref = (ops.bc_mask * (1 << np.arange(dim))).sum(axis=1)
    np.testing.assert_array_equal(st.bc_mask(), ref)
    st.set_boundary((1,))
    ops = mo.MAOperators(mesh, fixed_boundary_segments=[1])
    ref = (ops.bc_mask * (1 << np.arange(dim))).sum(axis=1)
    np.testing.assert_array_equal(st.bc_mask(), ref)
    from movement_b200 import _lib as L

    with pytest.raises(L.MovementB200Error):
        st.set_boundary((99,))


def _monitor_cases():
    from movement_b200 import _lib as L

    return [
        (2, L.MV_MON_BALL, (0.5, 0.5, 0.0, 0.4, 20.0, 200.0), mo.make_ball_monitor((0.5, 0.5), 0.4, 20.0, 200.0)),
        (2, L.MV_MON_BALL_RAW, (0.5, 0.5, 0.0, 0.15, 10.0, 200.0), mo.monitor_test_ring),
        (3, L.MV_MON_BALL_RAW, (0.5, 0.5, 0.5, 0.15, 10.0, 200.0), mo.monitor_test_ring),
        (2, L.MV_MON_GAUSS_HESS, (0.51, 0.65, 0.1, 5.0), mo.monitor_helmholtz_exact),
        (3, L.MV_MON_SINATAN3, (10.0,), mo.monitor_sinatan3),
        (2, L.MV_MON_CONST, (), mo.monitor_const),
    ]


@pytest.mark.parametrize("case", range(6))
def test_builtin_monitors(case):
    from movement_b200.core import MAState

    dim, kind, params, ref_fn = _monitor_cases()[case]
    mesh = _perturbed(_mesh(dim, 8 if dim == 2 else 5))
    dm = _dmesh(mesh)
    st = MAState(dm, mesh.coords)
    st.set_monitor(kind, params)
    out = dm.zeros(mesh.num_vertices)
    st.eval_monitor(st.x, out)
    ref = ref_fn(mesh.coords)
    np.testing.assert_allclose(out.cpu().numpy(), ref, rtol=1e-12, atol=1e-13)


# ---- tangling known-answer tests (reference test/test_tangling.py:39-95) -------------
def _find(coords, pt):
    return int(np.argmin(np.abs(coords - np.asarray(pt)).sum(axis=1)))


@pytest.mark.parametrize("shift,expected", [((0.2, 0.2), 1), ((0.5, 0.0), 2), ((0.5, 0.5), 3)])
def test_tangling_2d_kat(shift, expected):
    mesh = mo.unit_square_mesh(3)
    dm = _dmesh(mesh)
    x0 = torch.as_tensor(mesh.coords).cuda()
    detJ0 = dm.tangle_init(x0)
    assert dm.tangle_check(x0, detJ0) == 0
    c = mesh.coords.copy()
    c[_find(c, (1 / 3, 1 / 3))] += shift
    ratio = dm.empty(mesh.num_cells)
    n = dm.tangle_check(torch.as_tensor(c).cuda(), detJ0, ratio)
    assert n == expected
    ref = mo.OracleTanglingChecker(mesh.coords, mesh.cells, raise_error=False)
    np.testing.assert_array_equal(ratio.cpu().numpy(), ref.ratio(c))  # bit-exact


def test_tangling_3d_kat():
    mesh = mo.unit_cube_mesh(3)
    dm = _dmesh(mesh)
    x0 = torch.as_tensor(mesh.coords).cuda()
    detJ0 = dm.tangle_init(x0)
    c = mesh.coords.copy()
    c[_find(c, (0, 0, 1)), 2] -= 0.4
    assert dm.tangle_check(torch.as_tensor(c).cuda(), detJ0) == 2
    c = mesh.coords.copy()
    c[_find(c, (1 / 3, 1 / 3, 1 / 3)), 0] += 0.4
    assert dm.tangle_check(torch.as_tensor(c).cuda(), detJ0) == 6
    for axes, expected in (((0,), mesh.num_cells), ((0, 1), 0), ((0, 1, 2), mesh.num_cells)):
        c = mesh.coords.copy()
        for a in axes:
            c[:, a] = 1 - c[:, a]
        assert dm.tangle_check(torch.as_tensor(c).cuda(), detJ0) == expected


def test_tangling_1d_bit_exact():
    """A9 on interval meshes (movement/tangling.py:26-29 with d = 1): detJ = x1 - x0, ratio and count bit-exact."""
    from movement_b200.core import DeviceMesh

    rng = np.random.default_rng(12)
    n = 5000
    coords = np.sort(rng.random(n + 1))[:, None].copy()
    cells = np.stack([np.arange(n), np.arange(n) + 1], axis=1).astype(np.int32)
    dm = DeviceMesh(coords, cells, np.array([[0], [n]], dtype=np.int32), np.array([1, 2], dtype=np.int32))
    detJ0 = dm.tangle_init(torch.as_tensor(coords).cuda())
    moved = coords + 2e-4 * (rng.random((n + 1, 1)) - 0.5)
    ratio = dm.empty(n)
    cnt = dm.tangle_check(torch.as_tensor(moved).cuda(), detJ0, ratio)
    ref = mo.OracleTanglingChecker(coords, cells, raise_error=False).ratio(moved)
    assert np.array_equal(ratio.cpu().numpy(), ref)
    assert cnt == int((ref <= 0).sum()) and cnt > 0


@pytest.mark.parametrize("dim,n", [(2, 40), (3, 12)])
def test_tangling_bit_exact_random(dim, n):
    mesh = _mesh(dim, n)
    dm = _dmesh(mesh)
    rng = np.random.default_rng(7)
    c = mesh.coords + (0.6 / n) * (rng.random(mesh.coords.shape) - 0.5) * 2
    ref = mo.OracleTanglingChecker(mesh.coords, mesh.cells, raise_error=False)
    detJ0 = dm.tangle_init(torch.as_tensor(mesh.coords).cuda())
    np.testing.assert_array_equal(detJ0.cpu().numpy(), ref.detJ0)
    ratio = dm.empty(mesh.num_cells)
    nt = dm.tangle_check(torch.as_tensor(c).cuda(), detJ0, ratio)
    r = ref.ratio(c)
    assert nt == int((r <= 0).sum()) and nt > 0
    np.testing.assert_array_equal(ratio.cpu().numpy() <= 0, r <= 0)
    np.testing.assert_array_equal(ratio.cpu().numpy(), r)


@pytest.mark.parametrize("dim,n", [(2, 10), (3, 4)])
def test_cell_volume_stats(dim, n):
    mesh = _perturbed(_mesh(dim, n))
    dm = _dmesh(mesh)
    vol = dm.empty(mesh.num_cells)
    vmax, vmin, mean, cov = dm.cell_volumes(torch.as_tensor(mesh.coords).cuda(), vol)
    ref, _ = mo.cell_geometry(mesh.coords, mesh.cells)
    np.testing.assert_allclose(vol.cpu().numpy(), ref, rtol=1e-13)
    got = vol.cpu().numpy()
    assert vmax == got.max() and vmin == got.min()  # stats are exact functions of the stored field
    np.testing.assert_allclose([vmax, vmin, mean], [ref.max(), ref.min(), ref.mean()], rtol=1e-13)
    np.testing.assert_allclose(cov, mo.coefficient_of_variation(ref), rtol=1e-10)


# ---- whole relaxation iterations vs the oracle -------------------------------------------
def _relax_case(dim, n, kind, params, ref_monitor, niter):
    from movement_b200.core import MAState

    mesh = _mesh(dim, n)
    dm = _dmesh(mesh)
    st = MAState(dm, mesh.coords)
    st.set_boundary(())
    st.set_solver(rtol=1e-13)
    st.set_monitor(kind, params)
    trace, its = st.relax_iterate(0.1, niter)
    orc = mo.OracleRelaxation(mesh, ref_monitor, maxiter=niter + 1, rtol=0.0, dtol=1e300, raise_convergence_errors=False)
    try:
        orc.move()
    except Warning:
        pass
    return st, trace, its, orc


@pytest.mark.parametrize("case", [0, 1, 2, 3])
def test_relaxation_iterations_match_oracle(case):
    from movement_b200 import _lib as L

    dim, n, kind, params, fn = [
        (2, 20, L.MV_MON_BALL, (0.5, 0.5, 0.0, 0.4, 20.0, 200.0), mo.make_ball_monitor((0.5, 0.5), 0.4, 20.0, 200.0)),
        (2, 20, L.MV_MON_GAUSS_HESS, (0.51, 0.65, 0.1, 5.0), mo.monitor_helmholtz_exact),
        (3, 6, L.MV_MON_SINATAN3, (10.0,), mo.monitor_sinatan3),
        (3, 6, L.MV_MON_BALL_RAW, (0.5, 0.5, 0.5, 0.15, 10.0, 200.0), mo.monitor_test_ring),
    ][case]
    niter = 12
    st, trace, its, orc = _relax_case(dim, n, kind, params, fn, niter)
    for i in range(niter):
        _, vr, cov, res = orc.trace[i]
        np.testing.assert_allclose(trace[i, 1], res, rtol=1e-8, atol=1e-13)
        np.testing.assert_allclose(trace[i, 2] / trace[i, 3], vr, rtol=1e-9)
        np.testing.assert_allclose(trace[i, 5], cov, rtol=1e-8)
    # state after niter full iterations: oracle.move() ran niter+1 projections, compare phi/H
    np.testing.assert_allclose(st.phi_old.cpu().numpy() - st.phi_old.mean().item(), orc.phi_old - orc.phi_old.mean(), atol=1e-10)
    np.testing.assert_allclose(st.H_old.cpu().numpy(), orc.H_old, atol=1e-9)
    st.l2_project()
    st.update_coordinates()
    np.testing.assert_allclose(st.x.cpu().numpy(), orc.x, atol=1e-10)


def test_relaxation_bit_reproducible():
    from movement_b200 import _lib as L

    outs = []
    for _ in range(2):
        st, trace, its, _ = _relax_case(2, 16, L.MV_MON_GAUSS_HESS, (0.51, 0.65, 0.1, 5.0), mo.monitor_helmholtz_exact, 3)
        outs.append((st.phi_old.cpu().numpy().copy(), st.H_old.cpu().numpy().copy(), trace.copy()))
    np.testing.assert_array_equal(outs[0][0], outs[1][0])
    np.testing.assert_array_equal(outs[0][1], outs[1][1])
    np.testing.assert_array_equal(outs[0][2], outs[1][2])
