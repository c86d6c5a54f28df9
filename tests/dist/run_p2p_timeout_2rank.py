"""Launched by torchrun (2 ranks, NCCL, MV_P2P_TIMEOUT_S=2): a peer-memory halo exchange whose neighbour is late.

Exchange 1: both ranks on time -> ghosts = owners' values.  Exchange 2: rank 1 sleeps 6 s first -> rank 0's wait times out
after 2 s and the call returns MV_ERR_COMM (once); rank 1 finds rank 0's message waiting and succeeds.  Exchange 3, after a
barrier: both succeed with the right ghosts -- the time-out is an error of one call, not the end of the communicator
(ADVICE r2).  Prints 'DIST-P2P-TIMEOUT-OK' on rank 0."""
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

import movement_b200 as mv  # noqa: E402
from movement_b200 import _lib as L  # noqa: E402
from movement_b200 import core, parallel  # noqa: E402


def main():
    import ctypes as C

    parallel.init()
    rank = parallel.rank()
    lib = L.load()
    mesh = mv.UnitSquareMesh(24, 24)
    dm = core.device_mesh_from(mesh)
    info = (C.c_int64 * 4)()
    L.check(lib.mv_comm_info(info))
    gids = torch.as_tensor(np.asarray(mesh.part.global_vertex, dtype=np.float64)).cuda()
    nown = mesh.n_owned_vertices
    vec = torch.empty_like(gids)

    def exchange():
        vec[:nown] = gids[:nown]
        vec[nown:] = -1.0
        core.use_current_stream()
        rc = lib.mv_halo_exchange(dm.handle, 1, L.ptr(vec))
        torch.cuda.synchronize()
        return rc, bool(torch.equal(vec, gids))

    msgs, ok = [], True
    if not info[0]:
        msgs.append("peer-memory path not active on this box: nothing to time out")
    else:
        rc, good = exchange()
        ok = ok and rc == L.MV_OK and good
        msgs.append(f"rank {rank} exchange 1: rc {rc}, ghosts right {good}")
        torch.distributed.barrier()
        if rank == 1:
            time.sleep(6.0)
        t0 = time.time()
        rc, good = exchange()
        dt = time.time() - t0
        if rank == 0:
            ok = ok and rc == L.MV_ERR_COMM and "timed out" in L.last_error() and 1.5 < dt < 5.5
        else:
            ok = ok and rc == L.MV_OK and good
        msgs.append(f"rank {rank} exchange 2 (rank 1 six seconds late): rc {rc} after {dt:.2f} s, ghosts right {good}")
        torch.distributed.barrier()
        for k in (3, 4):
            rc, good = exchange()
            ok = ok and rc == L.MV_OK and good
            msgs.append(f"rank {rank} exchange {k}: rc {rc}, ghosts right {good}")
    t = torch.tensor([0.0 if ok else 1.0], device="cuda")
    torch.distributed.all_reduce(t)
    print("\n".join(msgs), flush=True)
    if rank == 0:
        print("DIST-P2P-TIMEOUT-OK" if t.item() == 0 else "DIST-P2P-TIMEOUT-FAIL", flush=True)
    parallel.finalize()
    sys.exit(0 if t.item() == 0 else 1)


if __name__ == "__main__":
    main()
