"""Launched by torchrun (world_size 2, NCCL): partitioned Monge-Ampere relaxation against the
CPU oracle on the global mesh.  Prints 'DIST-OK' on rank 0 when coordinates agree to 1e-10."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))

import movement_oracle as mo  # noqa: E402

import movement_b200 as mv  # noqa: E402
from movement_b200 import parallel  # noqa: E402


def main():
    dim = int(sys.argv[1]) if len(sys.argv) > 1 else 3
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 8
    parallel.init()
    rank, world = parallel.rank(), parallel.size()
    mesh = mv.UnitSquareMesh(n, n) if dim == 2 else mv.UnitCubeMesh(n, n, n)
    assert mesh.part is not None and mesh.part.size == world
    mon = mv.helmholtz_hessian_monitor() if dim == 2 else mv.sinatan3_hessian_monitor()
    ref_mon = mo.monitor_helmholtz_exact if dim == 2 else mo.monitor_sinatan3
    mover = mv.MongeAmpereMover(mesh, mon, method="relaxation", rtol=1e-4)
    nit = mover.move()
    x = mover.mesh.coordinates.tensor[: mesh.n_owned_vertices].cpu().numpy()
    gids = mesh.part.global_vertex[: mesh.n_owned_vertices]
    omesh = mo.unit_square_mesh(n) if dim == 2 else mo.unit_cube_mesh(n)
    o = mo.OracleRelaxation(omesh, ref_mon, rtol=1e-4)
    ref_nit = o.move()
    err = float(np.abs(x - o.x[gids]).max())
    t = torch.tensor([err], device="cuda", dtype=torch.float64)
    torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    # Laplacian smoother + tangling on the same partition
    mesh2 = mv.UnitSquareMesh(12, 12) if dim == 2 else mv.UnitCubeMesh(6, 6, 6)
    ls = mv.LaplacianSmoother(mesh2, 1.0)
    f = mv.Function(ls.coord_space).interpolate([0.05] + [0.0] * (dim - 1))
    bcs = [mv.DirichletBC(ls.coord_space, 0, [2]), mv.DirichletBC(ls.coord_space, f, [1])]
    ls.move(0.0, boundary_conditions=bcs)
    ok = nit == ref_nit and t.item() < 1e-10
    if rank == 0:
        print(f"iters {nit} vs oracle {ref_nit}; max coordinate error {t.item():.3e}; laplacian vr {ls.volume_ratio:.6f}")
        print("DIST-OK" if ok else "DIST-FAIL", flush=True)
    parallel.finalize()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
