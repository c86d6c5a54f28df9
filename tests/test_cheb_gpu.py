"""Fused Chebyshev mass solver (csrc/mv_cheb.cu): parity with the oracle's exact solve and with the
Jacobi-PCG it replaces, on every component count of the path (A1: d, A8: d*d)."""
import numpy as np
import pytest
import torch

import movement_oracle as mo

pytestmark = pytest.mark.gpu


def _mesh(dim, n, amp=0.35):
    mesh = mo.unit_square_mesh(n) if dim == 2 else mo.unit_cube_mesh(n)
    rng = np.random.default_rng(11)
    c = mesh.coords.copy()
    interior = np.all((c > 1e-12) & (c < 1 - 1e-12), axis=1)
    c[interior] += amp / n * (rng.random((interior.sum(), dim)) - 0.5)
    return mo.OMesh(c, mesh.cells, mesh.bfacets, mesh.btags)


def _dmesh(mesh):
    from movement_b200.core import DeviceMesh

    return DeviceMesh(mesh.coords, mesh.cells, mesh.bfacets, mesh.btags)


@pytest.fixture
def solver():
    from movement_b200 import _lib as L

    lib = L.load()

    def choose(kind):
        L.check(lib.mv_set_mass_solver(kind))

    yield choose
    L.check(lib.mv_set_mass_solver(L.MV_MASS_CHEBYSHEV))


@pytest.mark.parametrize("dim,n,ncomp", [(2, 24, 1), (2, 24, 2), (2, 24, 4), (3, 7, 1), (3, 7, 3), (3, 7, 9), (3, 12, 9)])
def test_chebyshev_matches_exact_solve_and_pcg(dim, n, ncomp, solver):
    from movement_b200 import _lib as L

    mesh = _mesh(dim, n)
    dm = _dmesh(mesh)
    ops = mo.MAOperators(mesh)
    V = mesh.num_vertices
    b = np.random.default_rng(5).standard_normal((V, ncomp)) * ops.ml[:, None]
    bt = torch.as_tensor(b).cuda()
    ref = ops.solve_mass(b)
    out = {}
    for kind in (L.MV_MASS_CHEBYSHEV, L.MV_MASS_CHEBYSHEV_FP64, L.MV_MASS_PCG):
        solver(kind)
        x = dm.zeros(V, ncomp) if ncomp > 1 else dm.zeros(V)
        it, rel = dm.pcg(L.MV_OP_MASS, bt if ncomp > 1 else bt[:, 0].contiguous(), x, rtol=1e-12)
        out[kind] = (it, x.cpu().numpy().reshape(V, ncomp))
        assert rel.max() <= 1e-12
        assert np.abs(out[kind][1] - ref).max() < 1e-10 * np.abs(ref).max()
    it_c, it_p = out[L.MV_MASS_CHEBYSHEV_FP64][0], out[L.MV_MASS_PCG][0]
    # a-priori bound: 2 rho^k <= 1e-12 with rho = (sqrt(kappa)-1)/(sqrt(kappa)+1), kappa = 2 + d
    kappa = 2.0 + dim
    rho = (np.sqrt(kappa) - 1) / (np.sqrt(kappa) + 1)
    assert it_c <= int(np.ceil(np.log(2e12 * np.sqrt(kappa)) / -np.log(rho))) + 2
    assert it_c < 2 * it_p
    # the oracle's restatement of the same iteration (same interval, same check schedule): same count
    x_o, it_o = ops.solve_mass_chebyshev(b, rtol=1e-12)
    assert abs(it_c - it_o) <= 1
    assert np.abs(out[L.MV_MASS_CHEBYSHEV_FP64][1] - x_o.reshape(V, ncomp)).max() < 1e-10 * np.abs(ref).max()
    assert np.abs(out[L.MV_MASS_CHEBYSHEV][1] - out[L.MV_MASS_PCG][1]).max() < 1e-10 * np.abs(ref).max()
    # the default solver: the same polynomial as single-precision sweeps between double-precision residual passes;
    # the oracle's restatement takes the same sweeps (3 residual passes from a cold start: two sweeps to 1e-12)
    x_m, it_m, passes = ops.solve_mass_chebyshev_mixed(b, rtol=1e-12)
    assert abs(out[L.MV_MASS_CHEBYSHEV][0] - it_m) <= 1 and passes == 3
    assert out[L.MV_MASS_CHEBYSHEV][0] <= it_c + 5
    assert np.abs(out[L.MV_MASS_CHEBYSHEV][1] - x_m.reshape(V, ncomp)).max() < 1e-10 * np.abs(ref).max()


@pytest.mark.parametrize("dim,n", [(2, 16), (3, 6)])
def test_chebyshev_dirichlet_mask(dim, n):
    from movement_b200 import _lib as L

    mesh = _mesh(dim, n)
    dm = _dmesh(mesh)
    ops = mo.MAOperators(mesh)
    V = mesh.num_vertices
    b = np.random.default_rng(6).standard_normal((V, dim))
    mask = (ops.bc_mask * (1 << np.arange(dim))).sum(axis=1).astype(np.uint16)
    x = dm.zeros(V, dim)
    it, rel = dm.pcg(L.MV_OP_MASS, torch.as_tensor(b).cuda(), x, mask=torch.as_tensor(mask.view(np.int16)).cuda(), rtol=1e-13)
    ref = np.stack([ops.solve_mass_bc(b[:, a], a) for a in range(dim)], axis=1)
    got = x.cpu().numpy()
    assert np.abs(got - ref).max() < 1e-10 * np.abs(ref).max()
    assert np.all(got[ops.bc_mask] == 0.0)


def test_chebyshev_warm_start_and_reproducibility():
    from movement_b200 import _lib as L

    mesh = _mesh(3, 8)
    dm = _dmesh(mesh)
    ops = mo.MAOperators(mesh)
    V = mesh.num_vertices
    b = np.random.default_rng(7).standard_normal((V, 9)) * ops.ml[:, None]
    bt = torch.as_tensor(b).cuda()
    x0 = dm.zeros(V, 9)
    it0, _ = dm.pcg(L.MV_OP_MASS, bt, x0, rtol=1e-12)
    x1 = dm.zeros(V, 9)
    it1, _ = dm.pcg(L.MV_OP_MASS, bt, x1, rtol=1e-12)
    assert it0 == it1 and torch.equal(x0, x1)            # deterministic reductions: identical bits
    # start from a guess that is already 1e-6 accurate: about half the iterations
    xw = x0 * (1.0 + 1e-6)
    itw, rel = dm.pcg(L.MV_OP_MASS, bt, xw, rtol=1e-12)
    assert itw < it0 - 5 and rel.max() <= 1e-12
    assert np.abs(xw.cpu().numpy() - x0.cpu().numpy()).max() < 1e-10 * np.abs(x0.cpu().numpy()).max()
    # converged input: zero iterations, the guess is returned untouched
    xc = x0.clone()
    itc, _ = dm.pcg(L.MV_OP_MASS, bt, xc, rtol=1e-9)
    assert itc == 0 and torch.equal(xc, x0)


def test_chebyshev_maxit_reports_not_converged():
    from movement_b200 import _lib as L

    mesh = _mesh(2, 12)
    dm = _dmesh(mesh)
    V = mesh.num_vertices
    b = torch.as_tensor(np.random.default_rng(8).standard_normal(V)).cuda()
    with pytest.raises(L.MovementB200Error, match="not converged in 4 iterations") as e:
        dm.pcg(L.MV_OP_MASS, b, dm.zeros(V), rtol=1e-12, maxit=4)
    assert e.value.status == L.MV_ERR_NOT_CONVERGED


@pytest.mark.parametrize("dim,n,ncomp", [(2, 96, 2), (2, 200, 4), (3, 64, 3), (3, 70, 9)])
def test_chebyshev_tiled_slice_schedule(dim, n, ncomp, solver, monkeypatch):
    """Meshes large enough for the lattice-tiled slice schedule (csrc/mv_sched.cu): the order in which the
    slices are visited must not change a single bit of the solution, only the reduction order of the
    measured residual; checked against the Jacobi-PCG, which does not use the schedule."""
    from movement_b200 import _lib as L
    from movement_b200.core import DeviceMesh
    from movement_b200.mesh import cube_arrays, square_arrays

    monkeypatch.setenv("MV_SCHED", "1")          # the identity schedule is the default (csrc/mv_sched.cu)
    arrays = square_arrays(n, n) if dim == 2 else cube_arrays(n, n, n)
    dm = DeviceMesh(*arrays)
    V = dm.Vown
    g = torch.Generator(device="cuda").manual_seed(3)
    b = torch.randn(V, ncomp, dtype=torch.float64, device="cuda", generator=g) / V
    assert dm.slice_schedule() == "none"
    x = torch.zeros_like(b)
    it, rel = dm.pcg(L.MV_OP_MASS, b, x, rtol=1e-12)
    assert dm.slice_schedule() == "tiled"
    assert rel.max() <= 1e-12
    solver(L.MV_MASS_PCG)
    y = torch.zeros_like(b)
    dm.pcg(L.MV_OP_MASS, b, y, rtol=1e-12)
    assert float((x - y).abs().max()) < 1e-10 * float(y.abs().max())
    # every row was visited: M x = b row by row
    r = dm.spmv(L.MV_OP_MASS, x) - b
    assert float(r.abs().max()) < 1e-10 * float(b.abs().max())
