"""CPU checks of the Galerkin initial guess of the Poisson solve (csrc/mv_ma.cu: k_hist_dots / k_hist_solve /
k_hist_guess, restated as oracle/movement_oracle.py: galerkin_guess; it prepares the solve of
movement/monge_ampere.py:441-448): energy optimality, exactness on the span, truncation when the history is
(numerically) dependent, and what it buys on a relaxation trajectory."""
import numpy as np
import scipy.sparse.linalg as spl

import movement_oracle as mo


def _poisson(n=10):
    ops = mo.MAOperators(mo.unit_cube_mesh(n))
    K = ops.K.tocsc()
    V = K.shape[0]
    # the constant nullspace is removed by pinning one vertex: an SPD system with the same structure
    keep = np.arange(1, V)
    return K[keep][:, keep].tocsc()


def test_guess_is_the_energy_projection_onto_the_history():
    K = _poisson()
    rng = np.random.default_rng(0)
    xs = [rng.standard_normal(K.shape[0]) for _ in range(3)]
    bs = [K @ x for x in xs]
    b = rng.standard_normal(K.shape[0])
    x0, used = mo.galerkin_guess(xs, bs, b)
    assert used == 3
    r = b - K @ x0
    # the residual is orthogonal to every history vector (Galerkin condition = minimal energy-norm error on the span)
    assert max(abs(float(np.dot(x, r))) for x in xs) < 1e-9 * np.linalg.norm(b) * max(np.linalg.norm(x) for x in xs)
    exact = spl.spsolve(K, b)
    e0 = exact - x0
    for x in xs:                           # no other point of the span is closer in the energy norm
        e1 = exact - (x0 + 1e-3 * x)
        assert float(e0 @ (K @ e0)) <= float(e1 @ (K @ e1)) + 1e-12
    # a right-hand side inside the span is reproduced exactly
    x1, _ = mo.galerkin_guess(xs, bs, 2.0 * bs[0] - 0.5 * bs[2])
    assert np.abs(x1 - (2.0 * xs[0] - 0.5 * xs[2])).max() < 1e-9


def test_dependent_history_is_truncated_not_amplified():
    K = _poisson(6)
    rng = np.random.default_rng(1)
    x = rng.standard_normal(K.shape[0])
    xs = [x, x * (1.0 + 1e-13), rng.standard_normal(K.shape[0])]
    bs = [K @ v for v in xs]
    b = K @ (0.3 * x)
    x0, used = mo.galerkin_guess(xs, bs, b)
    assert used == 1                       # the copy adds nothing: the factorisation stops there (newest-first order)
    assert np.abs(x0 - 0.3 * x).max() < 1e-9
    assert mo.galerkin_guess([], [], b)[1] == 0


def test_history_shortens_the_solves_of_a_relaxation():
    """Right-hand sides of consecutive pseudo-timesteps (a smoothly decaying sequence): the initial residual of the guess
    is orders of magnitude below that of a zero start, and depth 3 is better than depth 1."""
    K = _poisson(8)
    rng = np.random.default_rng(2)
    modes = [rng.standard_normal(K.shape[0]) for _ in range(4)]
    rates = [0.9, 0.8, 0.6, 0.3]
    rhs = lambda k: sum(r ** k * m for r, m in zip(rates, modes))   # noqa: E731
    hist_x, hist_b, red1, red3 = [], [], [], []
    for k in range(12):
        b = rhs(k)
        for depth, out in ((1, red1), (3, red3)):
            x0, _ = mo.galerkin_guess(hist_x[:depth], hist_b[:depth], b)
            out.append(np.linalg.norm(b - K @ x0) / np.linalg.norm(b))
        x = spl.spsolve(K, b)
        hist_x.insert(0, x)
        hist_b.insert(0, b)
    assert max(red1[4:]) < 0.2 and max(red3[4:]) < 0.02
    assert all(r3 < r1 for r1, r3 in zip(red1[4:], red3[4:]))
