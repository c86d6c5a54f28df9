"""Vertex-owner domain decomposition for the multi-GPU path.

replaces: the DMPlex partition + PyOP2 halo construction the reference inherits from Firedrake
(movement/mover.py:172 ``data_with_halos``, movement/tangling.py:45).

Owner computes: rank r owns a contiguous block of global vertex numbers; it holds every cell
that touches an owned vertex plus the 1-ring of ghost vertices those cells reference, so each
owned matrix row / assembled vertex value is complete locally and only *forward* (owned ->
ghost) halo exchange is ever needed.  Local numbering: owned vertices first (global order),
then ghosts grouped by owner rank (ascending) and sorted by global id -- each neighbour's
ghosts are one contiguous receive segment.  A cell is owned by the rank that owns its
smallest vertex; owned cells come first.  Pure numpy: runs (and is tested) without a GPU.
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np


def block_ranges(n, size):
    """Balanced contiguous blocks: rank r owns [starts[r], starts[r+1])."""
    base, rem = divmod(int(n), int(size))
    counts = np.full(size, base, dtype=np.int64)
    counts[:rem] += 1
    return np.concatenate([[0], np.cumsum(counts)])


@dataclass
class LocalPart:
    rank: int
    size: int
    coords: np.ndarray            # [Vloc,d]
    cells: np.ndarray             # [Cloc,d+1] local vertex numbers
    bfacets: np.ndarray           # [Floc,d]   local vertex numbers (facets touching an owned vertex)
    btags: np.ndarray
    n_owned_vertices: int
    n_owned_cells: int
    global_vertex: np.ndarray     # [Vloc] local -> global
    global_cell: np.ndarray       # [Cloc]
    global_num_vertices: int
    global_num_cells: int
    vertex_starts: np.ndarray     # [size+1] ownership ranges
    # halo plan
    neighbours: np.ndarray = field(default_factory=lambda: np.zeros(0, np.int32))
    send_ptr: np.ndarray = field(default_factory=lambda: np.zeros(1, np.int64))
    send_idx: np.ndarray = field(default_factory=lambda: np.zeros(0, np.int32))   # local owned vertices to send
    recv_ptr: np.ndarray = field(default_factory=lambda: np.zeros(1, np.int64))   # ghost segments, relative to n_owned


def owner_of(gv, starts):
    return np.searchsorted(starts, gv, side="right") - 1


def partition_mesh(coords, cells, bfacets, btags, rank, size, starts=None):
    """Extract rank ``rank``'s part of a global mesh.  Deterministic; every rank can compute
    every other rank's part, which is how the send lists are derived without communication."""
    coords = np.asarray(coords)
    cells = np.asarray(cells)
    V = coords.shape[0]
    starts = block_ranges(V, size) if starts is None else np.asarray(starts, dtype=np.int64)
    lo, hi = int(starts[rank]), int(starts[rank + 1])

    def local_cells(r):
        l, h = int(starts[r]), int(starts[r + 1])
        touch = ((cells >= l) & (cells < h)).any(axis=1)
        return np.nonzero(touch)[0]

    cidx = local_cells(rank)
    cmin = cells[cidx].min(axis=1)
    own_c = (cmin >= lo) & (cmin < hi)
    # cells whose smallest vertex is owned first; both groups keep global order
    cidx = np.concatenate([cidx[own_c], cidx[~own_c]])
    n_owned_cells = int(own_c.sum())
    used = np.unique(cells[cidx])
    ghosts = used[(used < lo) | (used >= hi)]
    gown = owner_of(ghosts, starts)
    order = np.lexsort((ghosts, gown))
    ghosts, gown = ghosts[order], gown[order]
    gvert = np.concatenate([np.arange(lo, hi, dtype=np.int64), ghosts.astype(np.int64)])
    # global -> local
    g2l = np.full(V, -1, dtype=np.int64)
    g2l[gvert] = np.arange(gvert.size)
    lcells = g2l[cells[cidx]].astype(np.int32)
    # boundary facets with at least one owned vertex (the owned vertex's boundary term needs them);
    # their other vertices are in the 1-ring by construction
    bfacets = np.asarray(bfacets).reshape(-1, max(coords.shape[1], 1))
    if bfacets.shape[0]:
        ftouch = ((bfacets >= lo) & (bfacets < hi)).any(axis=1)
        lbf = g2l[bfacets[ftouch]].astype(np.int32)
        lbt = np.asarray(btags)[ftouch].astype(np.int32)
        assert (lbf >= 0).all()
    else:
        lbf, lbt = np.zeros((0, coords.shape[1]), np.int32), np.zeros(0, np.int32)
    # halo plan: receive segments
    neigh = np.unique(gown).astype(np.int32)
    recv_ptr = np.concatenate([[0], np.cumsum([(gown == r).sum() for r in neigh])]).astype(np.int64)
    # send lists: neighbour r's ghosts owned by me = my owned vertices in cells that also touch a vertex
    # owned by r (exactly the cells local to both ranks); sorted by global id == r's ghost order.
    gcells = cells[cidx]
    send_lists = []
    for r in neigh.tolist():
        l, h = int(starts[r]), int(starts[r + 1])
        shared = gcells[((gcells >= l) & (gcells < h)).any(axis=1)]
        mine = np.unique(shared[(shared >= lo) & (shared < hi)])
        send_lists.append((mine - lo).astype(np.int32))
    send_ptr = np.concatenate([[0], np.cumsum([len(s) for s in send_lists])]).astype(np.int64)
    send_idx = np.concatenate(send_lists).astype(np.int32) if send_lists else np.zeros(0, np.int32)
    return LocalPart(rank=rank, size=size, coords=coords[gvert].copy(), cells=lcells, bfacets=lbf, btags=lbt,
                     n_owned_vertices=hi - lo, n_owned_cells=n_owned_cells, global_vertex=gvert,
                     global_cell=cidx.astype(np.int64), global_num_vertices=V, global_num_cells=int(cells.shape[0]),
                     vertex_starts=starts, neighbours=neigh, send_ptr=send_ptr, send_idx=send_idx, recv_ptr=recv_ptr)
