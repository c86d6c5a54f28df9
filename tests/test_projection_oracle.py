"""CPU pins of the oracle's restatement of the cross-mesh projection (movement/monitor.py:191-201; Firedrake's supermesh
projection is third-party and absent: PARITY UNPINNED beyond these properties): projecting onto the same mesh is the
identity, a globally linear field is reproduced on any moved mesh, a smooth field converges at second order."""
import numpy as np
import pytest

import movement_oracle as mo


def _moved(mesh, amp=0.3, seed=1):
    n = round(mesh.num_vertices ** (1.0 / mesh.dim)) - 1
    x = mesh.coords.copy()
    interior = np.all((x > 1e-12) & (x < 1 - 1e-12), axis=1)
    x[interior] += amp / n * (np.random.default_rng(seed).random((int(interior.sum()), mesh.dim)) - 0.5)
    return mo.OMesh(x, mesh.cells, mesh.bfacets, mesh.btags)


@pytest.mark.parametrize("mesh", [mo.unit_square_mesh(7), mo.unit_cube_mesh(3)], ids=["2d", "3d"])
def test_projection_identity_and_linear_fields(mesh):
    u = np.random.default_rng(0).random(mesh.num_vertices)
    assert np.abs(mo.project_p1(mesh, u, mesh) - u).max() < 1e-13
    tgt = _moved(mesh)
    w = np.arange(1, mesh.dim + 1, dtype=float)
    assert np.abs(mo.project_p1(mesh, 1 + mesh.coords @ w, tgt) - (1 + tgt.coords @ w)).max() < 1e-12


def test_projection_is_second_order_for_a_smooth_field():
    f = lambda c: np.sin(3 * c[:, 0]) * np.cos(2 * c[:, 1])  # noqa: E731
    err = []
    for n in (8, 16, 32):
        src = mo.unit_square_mesh(n)
        tgt = _moved(src)
        err.append(np.abs(mo.project_p1(src, f(src.coords), tgt) - f(tgt.coords)).max())
    assert err[0] < 6e-3 and err[1] < err[0] / 2 and err[2] < err[1] / 3


def test_point_evaluation_finds_the_containing_cell():
    mesh = _moved(mo.unit_cube_mesh(3))
    pts = np.random.default_rng(2).random((200, 3))
    w = np.array([2.0, -1.0, 0.5])
    vals, lmin = mo.p1_evaluate(mesh, mesh.coords @ w, pts)
    assert lmin.min() > -1e-12 and np.abs(vals - pts @ w).max() < 1e-13
