"""mv_cell_volumes_static -- the cell-volume diagnostic for callers that cannot allocate device memory.

Some users of this library call the diagnostics from places where an allocation is off the table: a rank of a multi-GPU
run while a neighbour's peer-memory kernel spins (cudaMalloc synchronises the device and would wait on it), a process
whose HBM already belongs to torch's caching allocator or to a full-size mover (the diagnostic must not be what runs out
of memory), a monitoring thread that may not block on the driver's allocator.  The entry point tested here needs no mesh
handle and keeps all its scratch in one fixed __device__ array inside the library.  It has no counterpart in the
reference; what it must do is agree with mv_cell_volumes, the allocating entry point the movers use.

Checked, always through the C ABI (ctypes):
  * volumes, max, min, mean and coefficient of variation equal mv_cell_volumes' on the same arrays, to the tolerance the
    other volume tests use (1e-13 on the field and its mean, 1e-10 on the coefficient of variation);
  * twice in a row per mesh, and a small mesh right after a large one: the scratch is shared by all calls and keeps the
    previous call's partial sums, so stale contents must not leak into the next answer;
  * sizes beyond the scratch, host pointers and bad dimensions are refused with a message, not truncated or staged.
The file sorts last on purpose (pytest -x stops at the first failure, so nothing should hide behind this one) and works
both alone in a fresh process -- the static call is the first thing the library does there -- and after the other files.
"""
import ctypes as C

import numpy as np
import pytest
import torch

import movement_oracle as mo

pytestmark = pytest.mark.gpu

MAX_CELLS = 1 << 27   # include/movement_b200.h: 32768 CTAs x 4096 cells


def _jiggled(dim, n):
    mesh = mo.unit_square_mesh(n) if dim == 2 else mo.unit_cube_mesh(n)
    rng = np.random.default_rng(3)
    c = mesh.coords.copy()
    interior = np.all((c > 1e-12) & (c < 1 - 1e-12), axis=1)
    c[interior] += (0.2 / n) * (rng.random((int(interior.sum()), dim)) - 0.5)
    return c, mesh


def _static(dim, coords_d, cells_d):
    """One call of the no-allocation entry point on torch-owned device arrays."""
    from movement_b200 import _lib as L
    from movement_b200.core import use_current_stream

    ncell = cells_d.shape[0]
    vol = torch.full((ncell,), -1.0, dtype=torch.float64, device="cuda")
    st = (C.c_double * 4)()
    use_current_stream()
    L.check(L.load().mv_cell_volumes_static(dim, ncell, L.ptr(coords_d), L.ptr(cells_d), L.ptr(vol), st))
    return vol, tuple(st[:])


def _allocating(mesh, coords_d):
    from movement_b200.core import DeviceMesh

    dm = DeviceMesh(mesh.coords, mesh.cells, mesh.bfacets, mesh.btags)
    vol = dm.empty(mesh.num_cells)
    return vol, dm.cell_volumes(coords_d, vol)


def _same(got, want):
    (gv, gs), (wv, ws) = got, want
    np.testing.assert_allclose(gv.cpu().numpy(), wv.cpu().numpy(), rtol=1e-13, atol=0)
    np.testing.assert_allclose(gs[:3], ws[:3], rtol=1e-13)
    np.testing.assert_allclose(gs[3], ws[3], rtol=1e-10)
    g = gv.cpu().numpy()
    assert gs[0] == g.max() and gs[1] == g.min()   # exact functions of the field that was written


# 2D n=64: 8192 cells = two full CTAs; 3D n=12: 10368 cells = two CTAs and a ragged third; the small ones: one CTA
@pytest.mark.parametrize("dim,n_big,n_small", [(2, 64, 5), (3, 12, 2)])
def test_static_volumes_equal_allocating_entry_point(dim, n_big, n_small):
    cb, big = _jiggled(dim, n_big)
    cs, small = _jiggled(dim, n_small)
    cb_d, cs_d = torch.as_tensor(cb).cuda(), torch.as_tensor(cs).cuda()
    cellb_d = torch.as_tensor(big.cells.astype(np.int32)).cuda()
    cells_d = torch.as_tensor(small.cells.astype(np.int32)).cuda()
    # the static calls come first: in a fresh process this is the library's very first work on the device
    first = _static(dim, cb_d, cellb_d)
    second = _static(dim, cb_d, cellb_d)
    after_big = _static(dim, cs_d, cells_d)      # scratch still holds the big mesh's partials
    again = _static(dim, cs_d, cells_d)
    want_big = _allocating(big, cb_d)
    want_small = _allocating(small, cs_d)
    _same(first, want_big)
    _same(second, want_big)
    _same(after_big, want_small)
    _same(again, want_small)
    # fixed fold order: a repeated call returns the same bits, not merely the same digits
    assert first[1] == second[1] and after_big[1] == again[1]
    assert torch.equal(first[0], second[0])
    # and against the CPU restatement of the reference's formula
    ref, _ = mo.cell_geometry(cb, big.cells)
    np.testing.assert_allclose(first[0].cpu().numpy(), ref, rtol=1e-13)
    np.testing.assert_allclose(first[1][3], mo.coefficient_of_variation(ref), rtol=1e-10)


def test_static_volumes_refuse_what_they_cannot_do():
    from movement_b200 import _lib as L

    lib = L.load()
    c, mesh = _jiggled(2, 4)
    c_d = torch.as_tensor(c).cuda()
    cells_d = torch.as_tensor(mesh.cells.astype(np.int32)).cuda()
    vol = torch.empty(mesh.num_cells, dtype=torch.float64, device="cuda")
    st = (C.c_double * 4)()
    # one cell more than the scratch can fold: refused before anything is launched, with the limit in the message
    rc = lib.mv_cell_volumes_static(2, MAX_CELLS + 1, L.ptr(c_d), L.ptr(cells_d), L.ptr(vol), st)
    assert rc == L.MV_ERR_UNSUPPORTED
    msg = L.last_error()
    assert str(MAX_CELLS) in msg and "mv_cell_volumes" in msg
    # host arrays would need a staging buffer, i.e. an allocation
    rc = lib.mv_cell_volumes_static(2, mesh.num_cells, L.ptr(c), L.ptr(cells_d), L.ptr(vol), st)
    assert rc == L.MV_ERR_BADARG and "device pointers" in L.last_error()
    for dim, ncell in ((0, 4), (4, 4), (2, 0)):
        assert lib.mv_cell_volumes_static(dim, ncell, L.ptr(c_d), L.ptr(cells_d), L.ptr(vol), st) == L.MV_ERR_BADARG
    # a refused call leaves the entry point usable
    got = _static(2, c_d, cells_d)
    _same(got, _allocating(mesh, c_d))
