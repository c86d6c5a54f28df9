"""Launched by torchrun (world_size >= 2, NCCL): the multigrid-preconditioned Poisson solve on a
partitioned mesh against the exact CPU solve on the global mesh.  Prints 'DIST-AMG-OK' on rank 0."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import movement_oracle as mo  # noqa: E402

import movement_b200 as mv  # noqa: E402
from movement_b200 import _lib as L  # noqa: E402
from movement_b200 import core, parallel  # noqa: E402


def main():
    dim = int(sys.argv[1]) if len(sys.argv) > 1 else 3
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 20
    parallel.init()
    rank, world = parallel.rank(), parallel.size()
    mesh = mv.UnitSquareMesh(n, n) if dim == 2 else mv.UnitCubeMesh(n, n, n)
    part = mesh.part
    dm = core.device_mesh_from(mesh)
    omesh = mo.unit_square_mesh(n) if dim == 2 else mo.unit_cube_mesh(n)
    ops = mo.MAOperators(omesh)
    xg = omesh.coords
    f = np.sin(5 * xg[:, 0]) * np.cos(3 * xg[:, 1]) + (xg[:, -1] - 0.3) ** 2
    bg = ops.ml * f
    exact = ops.solve_poisson_nullspace(bg)
    exact -= exact.mean()
    nown = mesh.n_owned_vertices
    gids = part.global_vertex
    ok = True
    msgs = []
    for kind in ("amg", "jacobi"):
        dm.set_preconditioner(kind)
        b = torch.as_tensor(bg[gids[:nown]]).cuda()
        x = dm.zeros(dm.V)
        it, rel = dm.pcg(L.MV_OP_STIFFNESS, b, x, rtol=1e-12, maxit=20000, project_mean=True)
        err = float(np.abs(x[:nown].cpu().numpy() - exact[gids[:nown]]).max()) / float(np.abs(exact).max())
        t = torch.tensor([err], device="cuda", dtype=torch.float64)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        msgs.append(f"{kind}: {it} iterations, relative error {t.item():.2e}")
        ok = ok and t.item() < 1e-9
        if kind == "amg":
            info = dm.amg_info(L.MV_OP_STIFFNESS)
            ok = ok and info is not None and it < 60
            msgs.append(f"hierarchy on rank {rank}: {info}")
    if rank == 0:
        print("\n".join(msgs))
        print("DIST-AMG-OK" if ok else "DIST-AMG-FAIL", flush=True)
    parallel.finalize()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
