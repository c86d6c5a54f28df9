"""SpringMover with the assembled spring blocks and the block-diagonal multigrid preconditioner (mv_movers.cu,
`spring_solve_amg`; replaces the dense solve at movement/spring.py:266) against the oracle's exact row-replaced solve
(oracle/movement_oracle.py OracleSpringMover, movement/spring.py:183-235,246-283) and against the matrix-free Jacobi-PCG
path it takes over from on large meshes."""
import ctypes as C
import math
import os

import numpy as np
import pytest
import torch

import movement_oracle as mo

pytestmark = pytest.mark.gpu


def _mv():
    import movement_b200 as mv

    mv.mover._print = mv.spring._print = lambda *a, **k: None
    return mv


def _info(mover):
    from movement_b200 import _lib as L

    info = (C.c_int64 * 8)()
    L.check(L.load().mv_spring_solver_info(mover._dm.handle, info))
    return list(info)


def _forced(mv, n, steps, amg_min, dt=0.1, amplitude=0.2):
    """The forced top boundary of demos/lineal_spring.py:99-147 (displacement read on the MOVED mesh)."""
    os.environ["MV_SPRING_AMG_MIN"] = str(amg_min)
    try:
        mesh = mv.UnitSquareMesh(n, n)
        mover = mv.SpringMover(mesh, dt)
        top = mv.Function(mover.coord_space)
        moving = mv.DirichletBC(mover.coord_space, top, 4)
        fixed = mv.DirichletBC(mover.coord_space, 0, [1, 2, 3])
        nodes = np.asarray(moving.nodes)

        def update(t):
            x0 = mover.mesh.coordinates.dat.data_ro[nodes, 0]
            top.dat.data[nodes, 1] = amplitude * math.sin(2 * math.pi * t) * np.sin(math.pi * x0)

        for k in range(1, steps + 1):
            mover.move(dt * k, update_boundary_displacement=update, boundary_conditions=(fixed, moving))
        return mesh, mover
    finally:
        os.environ.pop("MV_SPRING_AMG_MIN", None)


def _oracle_forced(n, steps, dt=0.1, amplitude=0.2):
    omesh = mo.unit_square_mesh(n)
    ref = mo.OracleSpringMover(omesh, dt)
    top = np.unique(omesh.bfacets[omesh.btags == 4])
    fixed = np.unique(omesh.bfacets[np.isin(omesh.btags, [1, 2, 3])])
    for k in range(1, steps + 1):
        t = dt * k
        vals = np.zeros((len(top), 2))
        vals[:, 1] = amplitude * math.sin(2 * math.pi * t) * np.sin(math.pi * ref.x[top, 0])
        # later boundary conditions override earlier ones at shared vertices (the corners belong to both sets)
        ref.move(np.concatenate([fixed, top]), np.concatenate([np.zeros((len(fixed), 2)), vals]))
    return ref


@pytest.mark.parametrize("n", [32, 48])   # (<= 640 vertices: one level, no hierarchy to test)
def test_spring_multigrid_path_matches_the_oracle(n):
    mv = _mv()
    steps = 4
    mesh, mover = _forced(mv, n, steps, amg_min=16)
    info = _info(mover)
    assert info[0] >= 2 and info[1] >= 2, info          # both diagonal blocks have a real hierarchy
    ref = _oracle_forced(n, steps)
    err = np.abs(mover.mesh.coordinates.dat.data_ro - ref.x).max()
    assert err < 1e-10, err
    # mesh-independent iteration counts (exact block solves give 25; one V-cycle per block stays within a small factor)
    assert max(mover.cg_iterations) <= 60, mover.cg_iterations
    # the hierarchies were built once and reused while the mesh moved
    assert info[4] == steps and info[2] == 0, info


def test_spring_multigrid_path_equals_jacobi_path():
    mv = _mv()
    n, steps = 96, 3
    _, a = _forced(mv, n, steps, amg_min=16)
    _, j = _forced(mv, n, steps, amg_min=10**9)
    assert _info(j)[0] == 0 and _info(a)[0] >= 2
    d = float((a.mesh.coordinates.tensor - j.mesh.coordinates.tensor).abs().max())
    assert d < 1e-10, d
    assert max(a.cg_iterations) * 4 < min(j.cg_iterations), (a.cg_iterations, j.cg_iterations)


def test_spring_multigrid_rebuilds_when_the_constrained_set_changes():
    mv = _mv()
    n = 32
    os.environ["MV_SPRING_AMG_MIN"] = "16"
    try:
        mesh = mv.UnitSquareMesh(n, n)
        mover = mv.SpringMover(mesh, 1.0)
        fA = mv.Function(mover.coord_space).interpolate([0.0, 0.01])
        bcA = [mv.DirichletBC(mover.coord_space, fA, 4), mv.DirichletBC(mover.coord_space, 0, [1, 2, 3])]
        bcB = [mv.DirichletBC(mover.coord_space, 0, [1, 2])]
        omesh = mo.unit_square_mesh(n)
        ref = mo.OracleSpringMover(omesh, 1.0)
        nA4 = np.unique(omesh.bfacets[omesh.btags == 4])
        nA0 = np.unique(omesh.bfacets[np.isin(omesh.btags, [1, 2, 3])])
        mover.move(0.0, boundary_conditions=bcA)
        ref.move(np.concatenate([nA4, nA0]), np.concatenate([np.tile([0.0, 0.01], (len(nA4), 1)), np.zeros((len(nA0), 2))]))
        assert _info(mover)[2] == 1
        assert np.abs(mover.mesh.coordinates.dat.data_ro - ref.x).max() < 1e-10
        mover.move(1.0, boundary_conditions=bcA)
        ref.move(np.concatenate([nA4, nA0]), np.concatenate([np.tile([0.0, 0.01], (len(nA4), 1)), np.zeros((len(nA0), 2))]))
        assert _info(mover)[2] == 0                      # same constrained set: hierarchies reused
        nB = np.unique(omesh.bfacets[np.isin(omesh.btags, [1, 2])])
        mover.move(2.0, boundary_conditions=bcB)
        ref.move(nB, np.zeros((len(nB), 2)))
        assert _info(mover)[2] == 1                      # different Dirichlet rows: rebuilt
        assert np.abs(mover.mesh.coordinates.dat.data_ro - ref.x).max() < 1e-9
    finally:
        os.environ.pop("MV_SPRING_AMG_MIN", None)
