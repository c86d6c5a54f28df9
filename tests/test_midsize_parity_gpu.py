"""Whole-mover parity at sizes where the BENCHMARKED solver stack is live: a multigrid hierarchy of >= 3 levels with
single-precision cycle values, the Galerkin history guess of the Poisson solve (active from relaxation step 3), the
multi-CTA / TMA-staged Chebyshev mass solves -- against golden vectors of the CPU oracle (tests/golden/midsize_*.npz,
written by tests/golden/make_midsize_golden.py; exact inner solves).  movement/monge_ampere.py:488-540.

Bars (north_star): coordinates within 1e-10, residual trace within 1e-8, and the Euclidean true residual of every
multigrid-preconditioned CG solve, recomputed at exit, within 10 x rtol (rtol = 1e-12).
"""
import ctypes as C
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _run(name):
    import movement_b200 as mv
    from movement_b200 import _lib as L

    g = np.load(os.path.join(GOLD, name + ".npz"))
    dim, n, iters = int(g["dim"]), int(g["n"]), int(g["iterations"])
    lib = L.load()
    L.check(lib.mv_true_residual_enable(1))
    try:
        mesh = mv.UnitSquareMesh(n, n) if dim == 2 else mv.UnitCubeMesh(n, n, n)
        monitor = mv.helmholtz_hessian_monitor() if dim == 2 else mv.sinatan3_hessian_monitor()
        with pytest.warns(UserWarning):
            mover = mv.MongeAmpereMover(mesh, monitor, method="relaxation", maxiter=iters, rtol=0.0, dtol=1e300,
                                        raise_convergence_errors=False)
        with pytest.raises(Warning, match=f"Solver failed to converge in {iters} iterations."):
            mover.move()
        mx, last, cnt = (C.c_double * 16)(), (C.c_double * 16)(), (C.c_int64 * 16)()
        L.check(lib.mv_true_residual_get(mx, last, cnt))
    finally:
        L.check(lib.mv_true_residual_enable(0))
    return g, mover, float(mx[1]), int(cnt[1])


@pytest.mark.parametrize("name", ["midsize_3d_n32", "midsize_3d_n48", "midsize_2d_n256"])
def test_relaxation_trajectory_matches_oracle_golden(name):
    g, mover, true_relres, audited = _run(name)
    iters = int(g["iterations"])
    # the stack under test really is the benchmarked one
    info = mover._dm.amg_info()
    assert info is not None and info["levels"] >= 3, info
    assert len(mover.cg_iterations) == iters - 1
    # trace: residual, theta, volume ratio, sigma/mu of every iteration
    tr = np.array(mover.trace)
    assert tr.shape[0] == iters
    assert np.abs(tr[:, 3] - g["trace"][:, 2]).max() < 1e-8
    assert np.abs(tr[:, 4] - g["theta"]).max() < 1e-10
    np.testing.assert_allclose(tr[:, 1], g["trace"][:, 0], rtol=1e-8)
    np.testing.assert_allclose(tr[:, 2], g["trace"][:, 1], rtol=1e-7, atol=1e-12)
    # coordinates after the last evaluated iteration
    x = mover.x.dat.data_ro
    assert np.abs(x[g["sample_index"]] - g["x_sample"]).max() < 1e-10
    assert np.abs(x.sum(axis=0) - g["x_sum"]).max() < 1e-10 * x.shape[0]
    assert np.abs((x ** 2).sum(axis=0) - g["x_sumsq"]).max() < 1e-10 * x.shape[0]
    assert abs(np.abs(x - mover.xi.dat.data_ro).max() - float(g["disp_max"])) < 1e-10
    phi = mover.phi_old.dat.data_ro
    assert np.abs((phi - phi.mean())[g["sample_index"]] - g["phi_sample"]).max() < 1e-10
    # every Poisson solve was audited: ||b - K x|| / ||b|| of the returned iterate
    assert audited == iters - 1
    assert true_relres < 1e-11, true_relres
