"""The C/OpenMP restatement (timed CPU baseline) against the numpy oracle (exact inner solves)."""
import numpy as np
import pytest

import movement_oracle as mo
import oracle_c as oc


@pytest.mark.parametrize("dim,n,kind,params,fn", [
    (2, 12, oc.MON_GAUSS_HESS, (0.51, 0.65, 0.1, 5.0), mo.monitor_helmholtz_exact),
    (3, 5, oc.MON_SINATAN3, (10.0,), mo.monitor_sinatan3),
    (3, 5, oc.MON_BALL_RAW, (0.5, 0.5, 0.5, 0.15, 10.0, 200.0), mo.monitor_test_ring),
])
def test_c_port_iterations_match_numpy_oracle(dim, n, kind, params, fn):
    mesh = mo.unit_square_mesh(n) if dim == 2 else mo.unit_cube_mesh(n)
    c = oc.CRelaxation(mesh.coords, mesh.cells, mesh.bfacets, mesh.btags, kind, params, inner_rtol=1e-13, nthreads=2)
    niter = 6
    traces = [c.step(update=True)[0] for _ in range(niter)]
    o = mo.OracleRelaxation(mesh, fn, maxiter=niter + 1, rtol=0.0, dtol=1e300, raise_convergence_errors=False)
    with pytest.raises(Warning):
        o.move()
    for i in range(niter):
        _, vr, cov, res = o.trace[i]
        np.testing.assert_allclose(traces[i][:3], [vr, cov, res], rtol=1e-8)
    c.step(update=False)
    assert np.abs(c.x - o.x).max() < 1e-10
    assert np.abs(c.get(2) - o.H_old).max() < 1e-9


def test_c_port_boundary_geometry_matches_numpy_oracle():
    mesh = mo.unit_cube_mesh(3)
    fcell, fn, fa = oc.boundary_facet_geometry(mesh.coords, mesh.cells, mesh.bfacets)
    rc, rn, ra = mo.boundary_facet_geometry(mesh)
    np.testing.assert_array_equal(fcell, rc)
    np.testing.assert_allclose(fn, rn, atol=1e-15)
    np.testing.assert_allclose(fa, ra, rtol=1e-14)


def test_c_port_full_move_iteration_count():
    mesh = mo.unit_square_mesh(10)
    c = oc.CRelaxation(mesh.coords, mesh.cells, mesh.bfacets, mesh.btags, oc.MON_BALL_RAW, (0.5, 0.5, 0.0, 0.15, 10.0, 200.0))
    assert c.move(rtol=1e-3) == 20   # SURVEY 4.3 regression value
