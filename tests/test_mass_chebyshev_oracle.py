"""CPU checks of the two facts the fused Chebyshev mass solver (csrc/mv_cheb.cu) rests on, and of its oracle
restatement (oracle/movement_oracle.py: MAOperators.solve_mass_chebyshev):
  * the spectrum of diag(M)^-1 M lies in [1/2, 1 + d/2] for P1 simplices on ANY mesh (Wathen 1987), also after
    Dirichlet rows/columns are removed (principal submatrix: Cauchy interlacing);
  * the iteration with host-constant coefficients converges to the exact (sparse LU) solve within the a-priori
    count, warm starts shorten it, constrained dofs keep their values."""
import numpy as np
import pytest
import scipy.sparse as sp

import movement_oracle as mo


def _mesh(dim, n, amp, seed=3, stretch=1.0):
    mesh = mo.unit_square_mesh(n) if dim == 2 else mo.unit_cube_mesh(n)
    rng = np.random.default_rng(seed)
    c = mesh.coords.copy()
    interior = np.all((c > 1e-12) & (c < 1 - 1e-12), axis=1)
    c[interior] += amp / n * (rng.random((interior.sum(), dim)) - 0.5)
    c[:, 0] *= stretch                      # anisotropic cells
    return mo.OMesh(c, mesh.cells, mesh.bfacets, mesh.btags)


@pytest.mark.parametrize("dim,n,amp,stretch", [(2, 12, 0.0, 1.0), (2, 12, 0.9, 1.0), (2, 10, 0.8, 25.0),
                                               (3, 5, 0.0, 1.0), (3, 5, 0.8, 1.0), (3, 4, 0.7, 40.0)])
def test_wathen_interval_of_jacobi_scaled_mass_matrix(dim, n, amp, stretch):
    ops = mo.MAOperators(_mesh(dim, n, amp, stretch=stretch))
    M = ops.M.toarray()
    d = np.sqrt(np.diag(M))
    ev = np.linalg.eigvalsh(M / d[:, None] / d[None, :])
    assert ev.min() >= 0.5 - 1e-12 and ev.max() <= 1.0 + 0.5 * dim + 1e-12
    # with the axis-aligned Dirichlet dofs of monge_ampere.py:258-272 removed, per component
    for a in range(dim):
        free = ~ops.bc_mask[:, a]
        Mf = M[np.ix_(free, free)]
        df = np.sqrt(np.diag(Mf))
        evf = np.linalg.eigvalsh(Mf / df[:, None] / df[None, :])
        assert evf.min() >= 0.5 - 1e-12 and evf.max() <= 1.0 + 0.5 * dim + 1e-12


@pytest.mark.parametrize("dim,n,ncomp", [(2, 16, 1), (2, 16, 4), (3, 6, 3), (3, 6, 9)])
def test_oracle_chebyshev_matches_exact_solve(dim, n, ncomp):
    ops = mo.MAOperators(_mesh(dim, n, 0.5))
    rng = np.random.default_rng(9)
    b = rng.standard_normal((ops.V, ncomp)) * ops.ml[:, None]
    ref = ops.solve_mass(b)
    x, it = ops.solve_mass_chebyshev(b, rtol=1e-12)
    assert np.abs(x - ref).max() < 1e-10 * np.abs(ref).max()
    kappa = 2.0 + dim
    rho = (np.sqrt(kappa) - 1) / (np.sqrt(kappa) + 1)
    assert it <= int(np.ceil(np.log(2e12 * np.sqrt(kappa)) / -np.log(rho))) + 2
    # warm start: fewer iterations, same answer; converged guess: zero iterations, untouched
    xw, itw = ops.solve_mass_chebyshev(b, x0=ref * (1 + 1e-6), rtol=1e-12)
    assert itw < it - 5 and np.abs(xw - ref).max() < 1e-10 * np.abs(ref).max()
    xc, itc = ops.solve_mass_chebyshev(b, x0=x, rtol=1e-9)
    assert itc == 0 and np.array_equal(xc, x)


def test_oracle_chebyshev_dirichlet_dofs():
    ops = mo.MAOperators(_mesh(3, 5, 0.4))
    rng = np.random.default_rng(10)
    b = rng.standard_normal((ops.V, 3))
    x, it = ops.solve_mass_chebyshev(b, fixed=ops.bc_mask, rtol=1e-13)
    ref = np.stack([ops.solve_mass_bc(b[:, a], a) for a in range(3)], axis=1)
    assert np.abs(x - ref).max() < 1e-10 * np.abs(ref).max()
    assert np.all(x[ops.bc_mask] == 0.0)
