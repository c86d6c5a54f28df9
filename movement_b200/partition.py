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
    method: str = "block"         # "block": owned vertices are global_vertex[:n_owned] = a contiguous range; "rcb": any set


def owner_of(gv, starts):
    return np.searchsorted(starts, gv, side="right") - 1


def rcb_order(coords, size):
    """Recursive coordinate bisection of the vertices into ``size`` parts of (almost) equal counts: the vertex set is cut
    perpendicular to its longest extent, proportionally to the number of parts on either side, until every set is one
    part.  Returns ``(perm, starts)``: part r consists of the vertices ``perm[starts[r]:starts[r+1]]`` (old numbers, ascending
    inside a part so that the generator's locality survives).  Deterministic (ties are broken by vertex number), so every
    rank computes the same decomposition without communication.  This is the general-mesh alternative to the
    contiguous index blocks, which are slabs only for meshes numbered plane by plane."""
    coords = np.asarray(coords, dtype=np.float64)
    V = coords.shape[0]
    parts = []

    def split(idx, p):
        if p == 1:
            parts.append(np.sort(idx))
            return
        p1 = p // 2
        k = (idx.size * p1) // p
        x = coords[idx]
        axis = int(np.argmax(x.max(axis=0) - x.min(axis=0))) if idx.size else 0
        order = np.lexsort((idx, x[:, axis]))
        split(idx[order[:k]], p1)
        split(idx[order[k:]], p - p1)

    split(np.arange(V, dtype=np.int64), int(size))
    starts = np.concatenate([[0], np.cumsum([q.size for q in parts])]).astype(np.int64)
    return np.concatenate(parts) if parts else np.zeros(0, np.int64), starts


def partition_mesh(coords, cells, bfacets, btags, rank, size, starts=None, method=None):
    """Extract rank ``rank``'s part of a global mesh.  Deterministic; every rank can compute
    every other rank's part, which is how the send lists are derived without communication.

    ``method``: ``"block"`` (default) -- contiguous blocks of the global vertex numbers; ``"rcb"`` -- recursive coordinate
    bisection (``rcb_order``) for meshes whose numbering carries no locality: the mesh is renumbered part by part, cut into
    those blocks, and ``global_vertex`` / ``global_cell`` of the result refer to the CALLER's numbering again.  The default
    can be switched with the environment variable ``MV_PARTITION``."""
    import os

    method = method or os.environ.get("MV_PARTITION", "block")
    if method == "rcb" and starts is None and size > 1:
        coords = np.asarray(coords)
        cells = np.asarray(cells)
        perm, rstarts = rcb_order(coords, size)
        inv = np.empty(coords.shape[0], dtype=np.int64)
        inv[perm] = np.arange(coords.shape[0])
        bf = np.asarray(bfacets)
        lp = partition_mesh(coords[perm], inv[cells], inv[bf] if bf.size else bf, btags, rank, size, starts=rstarts, method="block")
        lp.global_vertex = perm[lp.global_vertex]
        lp.method = "rcb"
        return lp
    if method not in ("block", "rcb"):
        raise ValueError(f"unknown partition method {method!r} (block, rcb)")
    coords = np.asarray(coords)
    cells = np.asarray(cells)
    V = coords.shape[0]
    starts = block_ranges(V, size) if starts is None else np.asarray(starts, dtype=np.int64)
    return _cut_block(lambda gv: coords[gv].copy(), coords.shape[1], V, cells, None, int(cells.shape[0]), bfacets, btags,
                      rank, size, starts)


def _cut_block(coords_of, dim, V, cells, cell_gid, num_global_cells, bfacets, btags, rank, size, starts):
    """The block partition proper.  ``cells`` / ``bfacets`` may be CANDIDATE subsets of the global arrays (any superset of
    the entities that touch an owned vertex, in global order; ``cell_gid`` = their global cell numbers, None = they are the
    global arrays), ``coords_of(global ids)`` returns coordinates -- so a generated mesh never has to exist as a whole
    (``cube_part`` / ``square_part`` below)."""
    cells = np.asarray(cells)
    lo, hi = int(starts[rank]), int(starts[rank + 1])
    touch = ((cells >= lo) & (cells < hi)).any(axis=1)
    cidx = np.nonzero(touch)[0]
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
    bfacets = np.asarray(bfacets).reshape(-1, max(dim, 1))
    if bfacets.shape[0]:
        ftouch = ((bfacets >= lo) & (bfacets < hi)).any(axis=1)
        lbf = g2l[bfacets[ftouch]].astype(np.int32)
        lbt = np.asarray(btags)[ftouch].astype(np.int32)
        assert (lbf >= 0).all()
    else:
        lbf, lbt = np.zeros((0, dim), np.int32), np.zeros(0, np.int32)
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
    global_cell = cidx.astype(np.int64) if cell_gid is None else np.asarray(cell_gid, dtype=np.int64)[cidx]
    return LocalPart(rank=rank, size=size, coords=coords_of(gvert), cells=lcells, bfacets=lbf, btags=lbt,
                     n_owned_vertices=hi - lo, n_owned_cells=n_owned_cells, global_vertex=gvert,
                     global_cell=global_cell, global_num_vertices=V, global_num_cells=int(num_global_cells),
                     vertex_starts=starts, neighbours=neigh, send_ptr=send_ptr, send_idx=send_idx, recv_ptr=recv_ptr)
