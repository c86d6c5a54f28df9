"""Host-side multi-GPU logic on the CPU: the vertex-owner partition and its halo plan, single
process and world_size-2 over gloo (the N>1 path of SURVEY 8e without a GPU)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from movement_b200.mesh import cube_arrays, square_arrays
from movement_b200.partition import block_ranges, partition_mesh, rcb_order

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("arrays", [cube_arrays(5, 4, 6), square_arrays(7, 9), cube_arrays(2, 2, 2)])
@pytest.mark.parametrize("size", [1, 2, 3, 8])
def test_partition_covers_mesh_and_halo_plan_is_consistent(arrays, size):
    c, ce, bf, bt = arrays
    if size > c.shape[0]:
        pytest.skip("more ranks than vertices")
    parts = [partition_mesh(c, ce, bf, bt, r, size) for r in range(size)]
    assert sum(p.n_owned_vertices for p in parts) == c.shape[0]
    assert sum(p.n_owned_cells for p in parts) == ce.shape[0]
    owned_cells = np.concatenate([p.global_cell[: p.n_owned_cells] for p in parts])
    assert np.array_equal(np.sort(owned_cells), np.arange(ce.shape[0]))        # every cell owned exactly once
    for p in parts:
        lo, hi = p.vertex_starts[p.rank], p.vertex_starts[p.rank + 1]
        assert np.array_equal(p.global_vertex[: p.n_owned_vertices], np.arange(lo, hi))
        # every cell touching an owned vertex is local, with its vertices renumbered consistently
        touch = ((ce >= lo) & (ce < hi)).any(axis=1)
        assert touch.sum() == p.cells.shape[0]
        assert np.array_equal(p.global_vertex[p.cells], ce[p.global_cell])
        assert np.array_equal(p.coords, c[p.global_vertex])
        # boundary facets: exactly those with an owned vertex
        ftouch = ((bf >= lo) & (bf < hi)).any(axis=1)
        assert ftouch.sum() == p.bfacets.shape[0]
        for k, q in enumerate(p.neighbours):
            sent = p.global_vertex[p.send_idx[p.send_ptr[k]:p.send_ptr[k + 1]]]
            Q = parts[q]
            kk = list(Q.neighbours).index(p.rank)
            recv = Q.global_vertex[Q.n_owned_vertices + Q.recv_ptr[kk]: Q.n_owned_vertices + Q.recv_ptr[kk + 1]]
            assert np.array_equal(sent, recv)
        assert p.n_owned_vertices + p.recv_ptr[-1] == p.coords.shape[0]


def _shuffled(arrays, seed=0):
    """The same mesh with a vertex numbering that carries no locality."""
    c, ce, bf, bt = arrays
    rng = np.random.default_rng(seed)
    new_of_old = rng.permutation(c.shape[0])
    old_of_new = np.argsort(new_of_old)
    return c[old_of_new], new_of_old[ce].astype(ce.dtype), new_of_old[bf].astype(bf.dtype), bt


@pytest.mark.parametrize("arrays", [cube_arrays(6, 5, 7), square_arrays(12, 9)])
@pytest.mark.parametrize("size", [2, 3, 5, 8])
def test_rcb_partition_of_an_unordered_mesh(arrays, size):
    """Recursive coordinate bisection (general meshes): same invariants as the block partition, in the caller's
    numbering, and far fewer ghosts than index blocks of a numbering without locality."""
    c, ce, bf, bt = _shuffled(arrays)
    V = c.shape[0]
    perm, starts = rcb_order(c, size)
    assert np.array_equal(np.sort(perm), np.arange(V)) and starts[0] == 0 and starts[-1] == V
    assert np.diff(starts).max() - np.diff(starts).min() <= 1 + size          # balanced
    parts = [partition_mesh(c, ce, bf, bt, r, size, method="rcb") for r in range(size)]
    owned = np.concatenate([p.global_vertex[: p.n_owned_vertices] for p in parts])
    assert np.array_equal(np.sort(owned), np.arange(V))                       # every vertex owned exactly once
    owned_cells = np.concatenate([p.global_cell[: p.n_owned_cells] for p in parts])
    assert np.array_equal(np.sort(owned_cells), np.arange(ce.shape[0]))       # every cell owned exactly once
    for p in parts:
        assert p.method == "rcb"
        mine = np.zeros(V, bool)
        mine[p.global_vertex[: p.n_owned_vertices]] = True
        assert np.array_equal(p.global_vertex[: p.n_owned_vertices], perm[starts[p.rank]:starts[p.rank + 1]])
        touch = mine[ce].any(axis=1)
        assert touch.sum() == p.cells.shape[0]                                # every cell touching an owned vertex is local
        assert np.array_equal(p.global_vertex[p.cells], ce[p.global_cell])
        assert np.array_equal(p.coords, c[p.global_vertex])
        assert mine[bf].any(axis=1).sum() == p.bfacets.shape[0]
        for k, q in enumerate(p.neighbours):
            sent = p.global_vertex[p.send_idx[p.send_ptr[k]:p.send_ptr[k + 1]]]
            Q = parts[q]
            kk = list(Q.neighbours).index(p.rank)
            recv = Q.global_vertex[Q.n_owned_vertices + Q.recv_ptr[kk]: Q.n_owned_vertices + Q.recv_ptr[kk + 1]]
            assert np.array_equal(sent, recv)
        assert p.n_owned_vertices + p.recv_ptr[-1] == p.coords.shape[0]
    ghosts_rcb = sum(int(p.recv_ptr[-1]) for p in parts)
    ghosts_block = sum(int(partition_mesh(c, ce, bf, bt, r, size).recv_ptr[-1]) for r in range(size))
    assert ghosts_rcb < ghosts_block                                          # index blocks of a shuffled numbering: almost everything is a ghost
    # on the generator's own numbering the parts are boxes: compact extents
    c0 = arrays[0]
    perm0, starts0 = rcb_order(c0, size)
    for r in range(size):
        x = c0[perm0[starts0[r]:starts0[r + 1]]]
        assert np.prod(x.max(axis=0) - x.min(axis=0) + 1e-12) <= 1.0 / size * 3.5


_FIELDS = ("coords", "cells", "bfacets", "btags", "global_vertex", "global_cell", "vertex_starts", "neighbours", "send_ptr",
           "send_idx", "recv_ptr")


@pytest.mark.parametrize("shape,size", [((5, 4, 6), 1), ((5, 4, 6), 2), ((5, 4, 6), 3), ((5, 4, 6), 7), ((3, 3, 2), 8),
                                        ((8, 8, 8), 8), ((2, 2, 12), 5), ((16, 16, 16), 4)])
def test_cube_slab_generated_without_the_global_mesh_equals_the_cut_of_the_global_mesh(shape, size):
    """`cube_part` creates only the hexahedron layers around a rank's vertex block; every array of the result must equal
    what `partition_mesh` extracts from the whole mesh (bench.py and UnitCubeMesh use it on several GPUs)."""
    from movement_b200.mesh import cube_part

    arrays = cube_arrays(*shape)
    for rank in range(size):
        ref = partition_mesh(*arrays, rank, size)
        got = cube_part(*shape, rank, size)
        for f in _FIELDS:
            a, b = getattr(ref, f), getattr(got, f)
            assert a.dtype == b.dtype and a.shape == b.shape and np.array_equal(a, b), (f, rank)
        for f in ("n_owned_vertices", "n_owned_cells", "global_num_vertices", "global_num_cells", "rank", "size"):
            assert getattr(ref, f) == getattr(got, f), (f, rank)


@pytest.mark.parametrize("shape,size", [((7, 9), 1), ((7, 9), 2), ((7, 9), 3), ((7, 9), 8), ((3, 20), 4), ((32, 32), 5)])
def test_square_slab_generated_without_the_global_mesh_equals_the_cut_of_the_global_mesh(shape, size):
    from movement_b200.mesh import square_part

    for L in ((1.0, 1.0), (2.0, 0.5)):
        arrays = square_arrays(*shape, *L)
        for rank in range(size):
            ref = partition_mesh(*arrays, rank, size)
            got = square_part(*shape, rank, size, *L)
            for f in _FIELDS:
                a, b = getattr(ref, f), getattr(got, f)
                assert a.dtype == b.dtype and a.shape == b.shape and np.array_equal(a, b), (f, rank)
            assert (ref.n_owned_vertices, ref.n_owned_cells, ref.global_num_cells) == (got.n_owned_vertices, got.n_owned_cells, got.global_num_cells)


def test_unknown_partition_method():
    with pytest.raises(ValueError, match="unknown partition method"):
        partition_mesh(*square_arrays(3, 3), 0, 2, method="metis")


def test_block_ranges():
    assert block_ranges(10, 3).tolist() == [0, 4, 7, 10]
    assert block_ranges(4, 4).tolist() == [0, 1, 2, 3, 4]


_WORKER = r'''
import os, sys
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from movement_b200 import parallel
from movement_b200.mesh import UnitCubeMesh, cube_arrays
assert parallel.init(backend="gloo")
rank, size = parallel.rank(), parallel.size()
mesh = UnitCubeMesh(4, 4, 4)                       # distributed because parallel is active
p = mesh.part
assert p is not None and p.size == size and mesh.n_owned_vertices == p.n_owned_vertices
# emulate mv_halo_exchange with gloo point-to-point: owned values = global id, ghosts must receive them
vec = torch.full((mesh.num_vertices(),), -1.0, dtype=torch.float64)
vec[: p.n_owned_vertices] = torch.from_numpy(p.global_vertex[: p.n_owned_vertices].astype(np.float64))
reqs, bufs = [], []
for k, q in enumerate(p.neighbours):
    send = vec[torch.from_numpy(p.send_idx[p.send_ptr[k]:p.send_ptr[k + 1]].astype(np.int64))].contiguous()
    recv = torch.empty(int(p.recv_ptr[k + 1] - p.recv_ptr[k]), dtype=torch.float64)
    reqs += [dist.isend(send, int(q)), dist.irecv(recv, int(q))]
    bufs.append((k, recv))
for r in reqs: r.wait()
for k, recv in bufs:
    vec[p.n_owned_vertices + p.recv_ptr[k]: p.n_owned_vertices + p.recv_ptr[k + 1]] = recv
assert np.array_equal(vec.numpy(), p.global_vertex.astype(np.float64)), "halo exchange plan wrong"
# global reductions used by the host logic
nv = parallel.allreduce_host(p.n_owned_vertices, "sum"); nc = parallel.allreduce_host(p.n_owned_cells, "sum")
assert nv == 125 and nc == 384
tags = [None] * size
dist.all_gather_object(tags, mesh.exterior_facets.unique_markers)
assert sorted(set(t for ts in tags for t in ts)) == [1, 2, 3, 4, 5, 6]
parallel.finalize()
print("GLOO-OK")
'''


@pytest.mark.parametrize("method", ["block", "rcb"])
def test_world_size_2_gloo(tmp_path, method):
    script = tmp_path / "worker.py"
    script.write_text(_WORKER)
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", MV_PARTITION=method)
    out = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                          "--master-addr", "127.0.0.1", "--master-port", "29533" if method == "block" else "29535", str(script), ROOT],
                         capture_output=True, text=True, timeout=300, env=env)
    assert out.returncode == 0, out.stdout[-3000:] + out.stderr[-3000:]
    assert out.stdout.count("GLOO-OK") == 2
