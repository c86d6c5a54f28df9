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


def test_c_port_boundary_geometry_with_vertex_numbers_beyond_the_int64_key_limit():
    """the facet key ((v0*V)+v1)*V+v2 overflowed int64 for V > 2.08 M vertices, i.e. above 126^3 cells, so
    the reference arm of bench.py died at its own default 128^3 sample.  Same small cube, its vertices scattered over
    3 M vertex numbers (3e6^3 = 2.7e19 > 2^63): adjacent cells, normals and areas must not change."""
    mesh = mo.unit_cube_mesh(3)
    want = oc.boundary_facet_geometry(mesh.coords, mesh.cells, mesh.bfacets)
    V_big = 3_000_000
    assert float(V_big) ** 3 > 2.0 ** 63
    where = np.sort(np.random.default_rng(5).choice(V_big, size=mesh.num_vertices, replace=False))
    where[-1] = V_big - 1
    coords = np.zeros((V_big, 3))
    coords[where] = mesh.coords
    got = oc.boundary_facet_geometry(coords, where[mesh.cells].astype(np.int32), where[mesh.bfacets].astype(np.int32))
    np.testing.assert_array_equal(got[0], want[0])
    np.testing.assert_array_equal(got[1], want[1])
    np.testing.assert_array_equal(got[2], want[2])


def test_reference_arm_of_bench_runs_end_to_end():
    """`bench.py --impl reference` (CPU only) prints one JSON line with the keys the driver reads."""
    import json
    import os
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--cpu-n", "6", "--steps", "2",
                          "--warmup", "1"], capture_output=True, text=True, timeout=600)
    assert out.returncode == 0, out.stderr[-2000:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["value"] > 0 and line["config"]["sample_n"] == 6
    assert line["e2e"]["value"] == line["value"] and line["cpu_baseline"]["kind"] == "port"
    assert "unavailable" not in line


def test_c_port_full_move_iteration_count():
    mesh = mo.unit_square_mesh(10)
    c = oc.CRelaxation(mesh.coords, mesh.cells, mesh.bfacets, mesh.btags, oc.MON_BALL_RAW, (0.5, 0.5, 0.0, 0.15, 10.0, 200.0))
    assert c.move(rtol=1e-3) == 20   # SURVEY 4.3 regression value
