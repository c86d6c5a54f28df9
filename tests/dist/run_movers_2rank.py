"""Launched by torchrun (world_size >= 2, NCCL): LaplacianSmoother, SpringMover and the quasi-Newton Monge-Ampere mover
on a partitioned mesh against the CPU oracle on the global mesh (movement/laplacian_smoothing.py:80-127,
movement/spring.py:246-283, movement/monge_ampere.py:729-750).  Prints 'DIST-MOVERS-OK' on rank 0."""
import math
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


def gmax(v):
    t = torch.tensor([float(v)], device="cuda", dtype=torch.float64)
    torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    return float(t.item())


def forced(kind, n, steps):
    """The forced top-boundary demo (demos/laplacian_smoothing.py:79-130, demos/lineal_spring.py:99-147)."""
    mesh = mv.UnitSquareMesh(n, n)
    mover = (mv.LaplacianSmoother if kind == "laplacian" else mv.SpringMover)(mesh, 0.1)
    top = mv.Function(mover.coord_space)
    moving = mv.DirichletBC(mover.coord_space, top, 4)
    fixed = mv.DirichletBC(mover.coord_space, 0, [1, 2, 3])
    src = mesh if kind == "laplacian" else mover.mesh

    def update(t):
        coord = src.coordinates.dat.data_ro
        data = top.dat.data
        for i in moving.nodes:
            data[i][1] = 0.2 * math.sin(2 * math.pi * t) * math.sin(math.pi * coord[i][0])

    kw = "update_boundary_velocity" if kind == "laplacian" else "update_boundary_displacement"
    omesh = mo.unit_square_mesh(n)
    ref = (mo.OracleLaplacianSmoother if kind == "laplacian" else mo.OracleSpringMover)(omesh, 0.1)
    top_nodes = np.unique(omesh.bfacets[omesh.btags == 4])
    fix_nodes = np.unique(omesh.bfacets[np.isin(omesh.btags, [1, 2, 3])])
    worst = 0.0
    for k in range(steps):
        t = 0.1 * k
        mover.move(t, boundary_conditions=(fixed, moving), **{kw: update})
        rc = omesh.coords if kind == "laplacian" else ref.x
        vals = np.zeros((len(top_nodes), 2))
        vals[:, 1] = 0.2 * math.sin(2 * math.pi * t) * np.sin(math.pi * rc[top_nodes, 0])
        # later conditions override earlier ones at shared vertices: fixed first, then the moving top
        ref.move(np.concatenate([fix_nodes, top_nodes]), np.concatenate([np.zeros((len(fix_nodes), 2)), vals]))
        no = mesh.n_owned_vertices
        x = mover.mesh.coordinates.tensor[:no].cpu().numpy()
        worst = max(worst, float(np.abs(x - ref.x[mesh.part.global_vertex[:no]]).max()))
    return gmax(worst)


def quasi_newton(n):
    mesh = mv.UnitSquareMesh(n, n)
    mover = mv.MongeAmpereMover(mesh, mv.helmholtz_hessian_monitor(), method="quasi_newton", rtol=1e-8)
    mover.ksp_rtol = 1e-9
    nit = mover.move()
    omesh = mo.unit_square_mesh(n)
    q = mo.OracleQuasiNewton(omesh, mo.monitor_helmholtz_exact, rtol=1e-8)
    ref_nit = q.move()
    mover.rtol = 2e-16; mover.snes_rtol = 0.0; mover.maxiter = 40
    mover.move()
    q2 = mo.OracleQuasiNewton(omesh, mo.monitor_helmholtz_exact, rtol=2e-16, maxiter=60, snes_rtol=0.0)
    q2.move()
    no = mesh.n_owned_vertices
    x = mover.mesh.coordinates.tensor[:no].cpu().numpy()
    return nit, ref_nit, gmax(np.abs(x - q2.x[mesh.part.global_vertex[:no]]).max())


def main():
    parallel.init()
    rank = parallel.rank()
    mv.mover._print = mv.laplacian_smoothing._print = mv.spring._print = mv.monge_ampere._print = lambda *a, **k: None
    e_lap = forced("laplacian", 12, 4)
    e_spr = forced("spring", 10, 4)
    # the same with the assembled blocks + block-diagonal multigrid preconditioner (default from 4096 vertices on)
    os.environ["MV_SPRING_AMG_MIN"] = "16"
    e_spr_mg = forced("spring", 40, 3)
    os.environ.pop("MV_SPRING_AMG_MIN")
    nit, ref_nit, e_qn = quasi_newton(20)
    ok = e_lap < 1e-10 and e_spr < 1e-10 and e_spr_mg < 1e-10 and abs(nit - ref_nit) <= 1 and e_qn < 1e-10
    if rank == 0:
        print(f"laplacian {e_lap:.3e}  spring {e_spr:.3e}  spring(multigrid) {e_spr_mg:.3e}  quasi-newton its {nit} vs {ref_nit}, x {e_qn:.3e}")
        print("DIST-MOVERS-OK" if ok else "DIST-MOVERS-FAIL", flush=True)
    parallel.finalize()
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
