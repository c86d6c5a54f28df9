"""CPU checks of the facts the block-diagonal multigrid preconditioner of the spring solve (csrc/mv_movers.cu,
`spring_solve_amg`; replaces the dense solve at movement/spring.py:266) rests on, with the oracle's stiffness matrix
(oracle/movement_oracle.py: spring_stiffness, movement/spring.py:148-181):
  * the diagonal blocks Kxx, Kyy are weighted graph Laplacians (symmetric, non-positive off-diagonals, zero row sums):
    the operators smoothed aggregation is built for;
  * K <= 2 diag(Kxx, Kyy) in the Loewner order, edge by edge (v v^T <= 2 diag(v^2) for a 2-vector v);
  * with the Dirichlet rows of the demos removed, CG preconditioned by EXACT block solves needs a number of iterations
    that does not grow with the mesh (25 on the unit-square meshes), while Jacobi grows like n."""
import numpy as np
import pytest
import scipy.sparse as sp
import scipy.sparse.linalg as spl

import movement_oracle as mo


def _stiffness(n, amp=0.0, seed=1):
    mesh = mo.unit_square_mesh(n)
    if amp:
        rng = np.random.default_rng(seed)
        c = mesh.coords.copy()
        interior = np.all((c > 1e-12) & (c < 1 - 1e-12), axis=1)
        c[interior] += amp / n * (rng.random((interior.sum(), 2)) - 0.5)
        mesh = mo.OMesh(c, mesh.cells, mesh.bfacets, mesh.btags)
    mover = mo.OracleSpringMover(mesh, 1.0)
    return mesh, mover.stiffness().tocsr()


@pytest.mark.parametrize("n,amp", [(8, 0.0), (8, 0.8), (13, 0.6)])
def test_diagonal_blocks_are_graph_laplacians_and_bound_the_operator(n, amp):
    mesh, K = _stiffness(n, amp)
    V = mesh.num_vertices
    x, y = np.arange(0, 2 * V, 2), np.arange(1, 2 * V, 2)
    Kxx, Kyy, Kxy = K[x][:, x].toarray(), K[y][:, y].toarray(), K[x][:, y].toarray()
    for B in (Kxx, Kyy):
        assert np.abs(B - B.T).max() < 1e-13
        off = B - np.diag(np.diag(B))
        assert off.max() <= 1e-14                                   # M-matrix sign pattern
        assert np.abs(B.sum(axis=1)).max() < 1e-11 * np.abs(B).max()   # constants in the kernel
    assert np.abs(Kxy - Kxy.T).max() < 1e-13
    # 2 diag(Kxx, Kyy) - K is positive semi-definite
    D = np.block([[Kxx, np.zeros_like(Kxx)], [np.zeros_like(Kyy), Kyy]])
    P = np.concatenate([x, y])
    Kp = K.toarray()[np.ix_(P, P)]
    ev = np.linalg.eigvalsh(2.0 * D - Kp)
    assert ev.min() > -1e-10 * np.abs(ev).max()


def _pcg_iterations(A, b, M):
    its = [0]

    def cb(_):
        its[0] += 1

    x, info = spl.cg(A, b, rtol=1e-12, M=M, callback=cb, maxiter=20000)
    assert info == 0
    return its[0]


def test_block_preconditioned_cg_is_mesh_independent():
    counts, jacobi = [], []
    for n in (16, 32, 64):
        mesh, K = _stiffness(n)
        V = mesh.num_vertices
        bn = np.unique(mesh.bfacets.ravel())
        free = np.ones(2 * V, bool)
        free[2 * bn] = False
        free[2 * bn + 1] = False
        idx = np.arange(2 * V)[free]
        isx = idx % 2 == 0
        perm = np.concatenate([np.nonzero(isx)[0], np.nonzero(~isx)[0]])
        Kff = K[free][:, free].tocsc()
        Kp = Kff[perm][:, perm].tocsc()
        nx = int(isx.sum())
        D = sp.block_diag([Kp[:nx, :nx], Kp[nx:, nx:]]).tocsc()
        lu = spl.splu(D)
        b = np.random.default_rng(0).standard_normal(Kp.shape[0])
        counts.append(_pcg_iterations(Kp, b, spl.LinearOperator(Kp.shape, lu.solve)))
        dinv = 1.0 / Kp.diagonal()
        jacobi.append(_pcg_iterations(Kp, b, spl.LinearOperator(Kp.shape, lambda r: dinv * r)))
    assert max(counts) <= 30 and max(counts) - min(counts) <= 3, counts       # 25 / 25 / 25 when this was written
    assert jacobi[2] > 1.7 * jacobi[1] > 1.7 * 1.7 * jacobi[0] * 0.9, jacobi  # O(n): 119 / 226 / 450
    assert jacobi[2] > 10 * counts[2]
