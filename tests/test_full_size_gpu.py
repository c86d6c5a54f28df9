"""BASELINE.json configs[1..4] at their FULL sizes, through the public API, checked by size-independent
properties (no oracle can run at these sizes):

  cfg2  LaplacianSmoother + MeshTanglingChecker, UnitSquareMesh 1024x1024   (demos/laplacian_smoothing.py forcing)
  cfg3  SpringMover (lineal), UnitSquareMesh 4096x4096                      (demos/lineal_spring.py forcing)
  cfg4  MongeAmpereMover relaxation, Helmholtz monitor, UnitSquareMesh 8192x8192; quasi-Newton at 4096x4096
        (FGMRES basis of the 8192^2 problem with the reference's restart does not fit one GPU; size stated)
  cfg5  3D relaxation 256^3: volume / boundary properties (the timed run is bench.py)

Properties (reference tests they generalise): rigid shift and forward+back (test/test_forced_movement.py:65-96),
linearity of the forced movers in the boundary data, discrete maximum principle for the Laplacian, fixed
boundaries stay exactly fixed, no tangling (test/test_tangling.py), volume and boundary preservation of the
Monge-Ampere movers (test/test_monge_ampere.py:262-355), bit-reproducibility.

MV_FULL_SIZE_DIV=k divides every mesh size by k (debugging on a small box)."""
import math
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
DIV = int(os.environ.get("MV_FULL_SIZE_DIV", "1"))


def _mv():
    import movement_b200 as mv

    mv.mover._print = lambda *a, **k: None
    mv.monge_ampere._print = mv.mover._print
    return mv


def _dev(mesh):
    """Moved coordinates as a device tensor [V, d] (no host round trip of 10^8 numbers)."""
    return mesh.coordinates.tensor


def _forced(mv, cls, mesh, scale, steps=1, dt=0.1, all_boundaries=None):
    """The demos' forcing (top boundary, 0.2 sin(2 pi t) sin(pi x), scaled) or a rigid shift of every boundary."""
    mover = cls(mesh, dt)
    if all_boundaries is not None:
        f = mv.Function(mover.coord_space).interpolate(list(all_boundaries))
        bcs = [mv.DirichletBC(mover.coord_space, f, [1, 2, 3, 4])]
        mover.move(0.0, boundary_conditions=bcs)
        return mover
    top = mv.Function(mover.coord_space)
    moving = mv.DirichletBC(mover.coord_space, top, 4)
    fixed = mv.DirichletBC(mover.coord_space, 0, [1, 2, 3])
    nodes = np.asarray(moving.nodes)
    x0 = mesh.coordinates.dat.data_ro[nodes, 0].copy()

    def update(t):
        data = top.dat.data
        data[nodes, 1] = scale * 0.2 * math.sin(2 * math.pi * t) * np.sin(math.pi * x0)

    kw = "update_boundary_velocity" if cls is mv.LaplacianSmoother else "update_boundary_displacement"
    for k in range(1, steps + 1):
        mover.move(dt * k, boundary_conditions=(fixed, moving), **{kw: update})
    return mover


@pytest.mark.parametrize("which,n", [("laplacian", 1024), ("spring", 4096)])
def test_forced_movers_full_size(which, n):
    mv = _mv()
    n //= DIV
    cls = mv.LaplacianSmoother if which == "laplacian" else mv.SpringMover
    mesh = mv.UnitSquareMesh(n, n)
    xi = _dev(mesh).clone()
    V = xi.shape[0]
    assert V == (n + 1) ** 2
    # rigid shift of the whole boundary: every vertex moves by exactly dt * v (harmonic / equilibrium extension of a constant)
    v = (0.3, -0.2)
    mover = _forced(mv, cls, mesh, 1.0, all_boundaries=v)
    shift = _dev(mover.mesh) - xi
    step = 0.1 * torch.tensor(v, dtype=torch.float64, device="cuda")   # x += v dt (laplacian_smoothing.py:120) / u dt (spring.py:268)
    assert float((shift - step).abs().max()) < 1e-9
    assert torch.equal(_dev(mesh), xi)                       # the caller's mesh is never touched (mover.py:49)
    del mover
    # demo forcing: one step at amplitude 1 and one at amplitude 2 -> linear in the data
    # (the spring stiffness follows the moving mesh, so its solve is not linear in the data: no amplitude-2 run)
    m1 = _forced(mv, cls, mesh, 1.0, steps=2 if which == "laplacian" else 5)
    if which == "spring":
        # block-diagonal multigrid preconditioner (mv_movers.cu): mesh-independent iteration counts, hierarchies reused
        print("spring cg iterations", m1.cg_iterations)
        assert DIV > 1 or max(m1.cg_iterations) <= 80, m1.cg_iterations
    d1 = _dev(m1.mesh) - xi
    scale = float(d1.abs().max())
    assert scale > 1e-3
    if which == "laplacian":                                 # K is assembled once on xi: exactly linear
        m2 = _forced(mv, cls, mesh, 2.0, steps=2)
        d2 = _dev(m2.mesh) - xi
        assert float((d2 - 2 * d1).abs().max()) < 1e-9 * max(scale, 1.0)
        del m2
    # fixed boundaries stay exactly where they were; the moving one only moves vertically
    c0 = xi
    left, right, bottom, top = c0[:, 0] == 0.0, c0[:, 0] == 1.0, c0[:, 1] == 0.0, c0[:, 1] == 1.0
    for msk in (left, right, bottom):
        assert float(d1[msk].abs().max()) == 0.0
    assert float(d1[top][:, 0].abs().max()) == 0.0
    # discrete maximum principle (Laplacian; M-matrix on this right-angled mesh): interior displacement bounded by the data
    if which == "laplacian":
        assert float(d1[:, 1].abs().max()) <= float(d1[top][:, 1].abs().max()) * (1 + 1e-12)
    # no tangling, through the checker the mover owns
    assert m1.tangling_checker.check() == 0
    ratio = m1.tangling_checker.jacobian_determinant_ratio.tensor
    assert ratio.shape[0] == 2 * n * n and float(ratio.min()) > 0.0


def _cell_volumes(mover):
    vol = torch.empty(mover._dm.C, dtype=torch.float64, device="cuda")
    mover._dm.cell_volumes(mover.mesh.coordinates.tensor, vol)
    return vol[: mover._dm.Cown]


def _run(mover):
    """move() with maxiter reached: the reference raises a Warning when raise_convergence_errors=False (mover.py:90-100)."""
    try:
        mover.move()
    except Warning:
        pass
    mover.to_physical_coordinates()


def _ma_properties(mover, mesh, dim):
    xi = _dev(mesh)
    x = _dev(mover.mesh)
    assert x.shape == xi.shape and bool(torch.isfinite(x).all())
    disp = x - xi
    assert float(disp.abs().max()) > 1e-6
    # boundary preservation: the normal component of every boundary vertex is untouched (monge_ampere.py:258-272)
    for a in range(dim):
        for val in (0.0, 1.0):
            msk = xi[:, a] == val
            assert float(disp[msk][:, a].abs().max()) == 0.0
    # volume preservation: the moved mesh still tiles the unit domain, and no cell is inverted
    assert mover.tangling_checker.check() == 0
    vol = _cell_volumes(mover)
    assert abs(float(vol.sum()) - 1.0) < 1e-11
    assert float(vol.min()) > 0.0
    assert np.isfinite(mover.relative_l2_residual) and mover.relative_l2_residual > 0.0


def test_monge_ampere_relaxation_2d_full_size():
    mv = _mv()
    n = 8192 // DIV
    mesh = mv.UnitSquareMesh(n, n)
    xs, res = [], []
    for rep in range(2):
        mover = mv.MongeAmpereMover(mesh, mv.helmholtz_hessian_monitor(), method="relaxation", rtol=1e-8, maxiter=3,
                                    raise_convergence_errors=False)
        _run(mover)
        if rep == 0:
            _ma_properties(mover, mesh, 2)
        xs.append(_dev(mover.mesh).clone())
        res.append(mover.relative_l2_residual)
        del mover
        torch.cuda.empty_cache()
    assert torch.equal(xs[0], xs[1]) and res[0] == res[1]    # deterministic reductions: identical bits run to run


def test_monge_ampere_quasi_newton_2d_large():
    mv = _mv()
    n = 4096 // DIV
    mesh = mv.UnitSquareMesh(n, n)
    mover = mv.MongeAmpereMover(mesh, mv.helmholtz_hessian_monitor(), method="quasi_newton", rtol=1e-8, maxiter=2,
                                raise_convergence_errors=False)
    mover.ksp_gmres_restart = 10
    _run(mover)
    _ma_properties(mover, mesh, 2)


def test_monge_ampere_relaxation_3d_full_size():
    mv = _mv()
    n = 256 // DIV
    mesh = mv.UnitCubeMesh(n, n, n)
    mover = mv.MongeAmpereMover(mesh, mv.sinatan3_hessian_monitor(), method="relaxation", rtol=1e-8, maxiter=3,
                                raise_convergence_errors=False)
    _run(mover)
    _ma_properties(mover, mesh, 3)
