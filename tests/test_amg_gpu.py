"""Multigrid-preconditioned CG (mv_amg*.cu) against the exact CPU solves of the oracle and against
the Jacobi-preconditioned path: same solution, an order of magnitude fewer iterations, and
bit-reproducible run to run.  The hierarchy itself is compared with oracle/amg_prototype.py."""
import numpy as np
import pytest
import scipy.sparse as sp
import scipy.sparse.linalg as spla
import torch

import movement_oracle as mo

pytestmark = pytest.mark.gpu


def _mesh(dim, n):
    return mo.unit_square_mesh(n) if dim == 2 else mo.unit_cube_mesh(n)


def _dmesh(mesh):
    from movement_b200.core import DeviceMesh

    return DeviceMesh(mesh.coords, mesh.cells, mesh.bfacets, mesh.btags)


def _perturbed(mesh, amp=0.25):
    rng = np.random.default_rng(3)
    c = mesh.coords.copy()
    h = 1.0 / round(mesh.num_vertices ** (1.0 / mesh.dim) - 1)
    interior = np.all((c > 1e-12) & (c < 1 - 1e-12), axis=1)
    c[interior] += amp * h * (rng.random((interior.sum(), mesh.dim)) - 0.5)
    return mo.OMesh(c, mesh.cells, mesh.bfacets, mesh.btags)


@pytest.mark.parametrize("dim,n,perturb", [(2, 48, False), (2, 40, True), (3, 12, False), (3, 16, True)])
def test_poisson_nullspace_amg_vs_exact(dim, n, perturb):
    """K u = b with constant nullspace (movement/monge_ampere.py:441-448)."""
    from movement_b200 import _lib as L

    mesh = _mesh(dim, n)
    if perturb:
        mesh = _perturbed(mesh)
    ops = mo.MAOperators(mesh)
    rng = np.random.default_rng(0)
    b = rng.standard_normal(mesh.num_vertices) * ops.ml
    exact = ops.solve_poisson_nullspace(b)
    exact -= exact.mean()
    dm = _dmesh(mesh)
    out = {}
    for kind in ("amg", "jacobi"):
        dm.set_preconditioner(kind)
        x = dm.zeros(mesh.num_vertices)
        it, rel = dm.pcg(L.MV_OP_STIFFNESS, torch.as_tensor(b).cuda(), x, rtol=1e-13, maxit=20000, project_mean=True)
        out[kind] = (it, x.cpu().numpy())
        scale = np.abs(exact).max()
        assert np.abs(out[kind][1] - exact).max() < 2e-10 * scale, (kind, it)
    info = dm.amg_info(L.MV_OP_STIFFNESS)
    assert info is not None and info["levels"] >= 2 and info["rows"][0] == mesh.num_vertices
    assert info["dense_coarse_solve"]
    assert out["amg"][0] < 45 and out["amg"][0] * 3 < out["jacobi"][0], (out["amg"][0], out["jacobi"][0], info)
    # operator complexity of the hierarchy stays modest
    assert sum(info["nnz"]) < 2.0 * info["nnz"][0], info


@pytest.mark.parametrize("dim,n,ncomp", [(2, 40, 2), (3, 12, 3), (2, 33, 1)])
def test_dirichlet_amg_vs_exact(dim, n, ncomp):
    """K_ff v_f = -K_fb g_b, all components constrained on the whole boundary
    (movement/laplacian_smoothing.py:55-77)."""
    from movement_b200 import _lib as L

    mesh = _perturbed(_mesh(dim, n))
    V = mesh.num_vertices
    ops = mo.MAOperators(mesh)
    bnodes = np.unique(mesh.bfacets)
    fixed = np.zeros(V, dtype=bool)
    fixed[bnodes] = True
    rng = np.random.default_rng(1)
    g = np.zeros((V, ncomp))
    g[fixed] = rng.standard_normal((fixed.sum(), ncomp))
    K = ops.K.tocsr()
    free = ~fixed
    Kff = K[free][:, free].tocsc()
    exact = g.copy()
    exact[free] = spla.splu(Kff).solve(-(K[free][:, fixed] @ g[fixed]))
    dm = _dmesh(mesh)
    mask = torch.as_tensor(np.where(fixed, (1 << ncomp) - 1, 0).astype(np.uint16).view(np.int16)).cuda()
    its = {}
    for kind in ("amg", "jacobi"):
        dm.set_preconditioner(kind)
        x = torch.as_tensor(g.copy()).cuda()
        if ncomp == 1:
            x = x.reshape(V).contiguous()
        it, _ = dm.pcg(L.MV_OP_STIFFNESS, None, x, mask=mask, rtol=1e-13, maxit=20000)
        its[kind] = it
        assert np.abs(x.cpu().numpy().reshape(V, ncomp) - exact).max() < 1e-10 * np.abs(exact).max(), kind
    assert its["amg"] < 40 and its["amg"] * 2 < its["jacobi"], its


def test_amg_bit_reproducible():
    from movement_b200 import _lib as L

    mesh = mo.unit_cube_mesh(10)
    ops = mo.MAOperators(mesh)
    b = torch.as_tensor(np.random.default_rng(2).standard_normal(mesh.num_vertices) * ops.ml).cuda()
    sols = []
    for _ in range(2):
        dm = _dmesh(mesh)            # fresh hierarchy each time
        x = dm.zeros(mesh.num_vertices)
        it, _ = dm.pcg(L.MV_OP_STIFFNESS, b, x, rtol=1e-12, project_mean=True)
        sols.append((it, x.cpu().numpy()))
    assert sols[0][0] == sols[1][0]
    assert np.array_equal(sols[0][1], sols[1][1])


def test_amg_mass_plus_stiffness():
    """(M + K) y = r: the phi block of the quasi-Newton fieldsplit (movement/solver_parameters.py:26-33)."""
    from movement_b200 import _lib as L

    mesh = _perturbed(mo.unit_square_mesh(36))
    ops = mo.MAOperators(mesh)
    A = (ops.M + ops.K).tocsc()
    b = np.random.default_rng(5).standard_normal(mesh.num_vertices) * ops.ml
    exact = spla.splu(A).solve(b)
    dm = _dmesh(mesh)
    x = dm.zeros(mesh.num_vertices)
    it, _ = dm.pcg(L.MV_OP_MASS_PLUS_STIFFNESS, torch.as_tensor(b).cuda(), x, rtol=1e-13)
    assert np.abs(x.cpu().numpy() - exact).max() < 1e-10 * np.abs(exact).max()
    info = dm.amg_info(L.MV_OP_MASS_PLUS_STIFFNESS)
    assert info is not None and it < 40, (it, info)


def test_amg_hierarchy_matches_prototype():
    """Aggregation is deterministic and equals the numpy statement of the algorithm: same coarse sizes."""
    import amg_prototype as ap
    from movement_b200 import _lib as L

    mesh = mo.unit_cube_mesh(14)
    ops = mo.MAOperators(mesh)
    dm = _dmesh(mesh)
    x = dm.zeros(mesh.num_vertices)
    b = torch.as_tensor(ops.ml * np.sin(7 * mesh.coords[:, 0])).cuda()
    dm.pcg(L.MV_OP_STIFFNESS, b, x, rtol=1e-10, project_mean=True)
    info = dm.amg_info()
    K = dm.to_scipy(L.MV_GET_STIFF).tocsr()      # the device's own bits, so that ties break identically
    lv = ap.build_hierarchy(K, theta=1e-8, max_coarse=640)
    assert [l.A.shape[0] for l in lv] == info["rows"], (info, [l.A.shape[0] for l in lv])
    assert abs(lv[0].rho - info["rho"][0]) < 1e-9 * info["rho"][0]
    # prolongator, restriction and Galerkin operator, entry by entry
    nc = info["rows"][1]
    P = dm.amg_matrix(0, "P", ncols=nc)
    R = dm.amg_matrix(0, "R", ncols=mesh.num_vertices)
    Ac = dm.amg_matrix(1, "A", ncols=nc)
    assert abs(P - lv[0].P).max() < 1e-13
    assert abs(R - P.T).max() == 0.0
    ref = lv[1].A
    assert abs(Ac - ref).max() < 1e-12 * abs(ref).max()
    assert abs(Ac - Ac.T).max() < 1e-13 * abs(ref).max()
    for l, Lv in enumerate(lv):
        assert abs(Lv.rho - info["rho"][l]) < 1e-6 * info["rho"][l], (l, Lv.rho, info["rho"][l])
