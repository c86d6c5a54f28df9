#!/usr/bin/env python
"""A CONVERGED Monge-Ampere relaxation through the public API: demos/monge_ampere_3d.py:74-78 (rtol = 1e-8,
mover.move()) on UnitCubeMesh(n,n,n), 1 GPU or partitioned under torchrun.

  python profiles/tools/converged_run.py --n 256 --out gpurun_out/converged_n256_N1.json [--record rec.npz]
  python -m torch.distributed.run --nproc-per-node 8 ... profiles/tools/converged_run.py --size 256 --compare rec.npz

Prints / stores: iteration count, final residual, residual trace, tangled cells of the PHYSICAL mesh at convergence,
wall time of move(), converged vertex-iterations/s, inner iteration totals, checksums of the moved coordinates; with
--record (one GPU) the coordinates at every `stride`-th vertex are stored, with --compare the run is checked
against such a record (same iteration count, max |x_N - x_1| over the sampled vertices).
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", "--size", dest="n", type=int, default=256, help="--size under torchrun (its parser finds --n ambiguous)")
    ap.add_argument("--dim", type=int, default=3)
    ap.add_argument("--rtol", type=float, default=1e-8)
    ap.add_argument("--maxiter", type=int, default=1000)
    ap.add_argument("--out", default=None)
    ap.add_argument("--record", default=None)
    ap.add_argument("--compare", default=None)
    ap.add_argument("--quiet", action="store_true")
    a = ap.parse_args()

    import numpy as np
    import torch
    import torch.distributed as dist

    import movement_b200 as mv
    from movement_b200 import parallel

    parallel.init()
    rank, world = parallel.rank(), parallel.size()
    n = a.n
    t0 = time.perf_counter()
    mesh = mv.UnitCubeMesh(n, n, n) if a.dim == 3 else mv.UnitSquareMesh(n, n)
    t_mesh = time.perf_counter() - t0
    monitor = mv.sinatan3_hessian_monitor() if a.dim == 3 else mv.helmholtz_hessian_monitor()
    if a.quiet:
        mv.monge_ampere._print = lambda *x, **k: None
    mover = mv.MongeAmpereMover(mesh, monitor, method="relaxation", rtol=a.rtol, maxiter=a.maxiter)
    torch.cuda.synchronize()
    t_setup = time.perf_counter() - t0 - t_mesh
    if world > 1:
        dist.barrier()
    t1 = time.perf_counter()
    its = mover.move()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    wall = time.perf_counter() - t1
    # tangling of the physical mesh at convergence (mesh.coordinates == x after move())
    tangled = mover.tangling_checker.check()
    V = mover._dm.global_vertices
    Vown = mover._dm.Vown
    x = mover.x.tensor[:Vown]
    sums = torch.cat([x.sum(dim=0), (x * x).sum(dim=0)])
    if world > 1:
        dist.all_reduce(sums)
    lp = mesh.part
    gv = lp.global_vertex[:Vown] if lp is not None else np.arange(V)
    stride = max(1, V // 16384)
    sel = np.nonzero(gv % stride == 0)[0]
    xs = x[torch.as_tensor(sel, device=x.device)].cpu().numpy()
    cmp_ = None
    if a.record and world == 1:
        np.savez_compressed(a.record, n=n, iterations=its, stride=stride, x_sample=xs,
                            residual=np.array([t[3] for t in mover.trace]))
    if a.compare and os.path.exists(a.compare):
        g = np.load(a.compare)
        dx = torch.tensor([float(np.abs(xs - g["x_sample"][gv[sel] // stride]).max()) if len(sel) else 0.0], device="cuda")
        if world > 1:
            dist.all_reduce(dx, op=dist.ReduceOp.MAX)
        m = min(len(mover.trace), len(g["residual"]))
        cmp_ = {"record": os.path.basename(a.compare), "iterations_record": int(g["iterations"]), "iterations_here": int(its),
                "max_abs_x_diff_sampled": float(dx.item()), "vertices_sampled": int(len(g["x_sample"])),
                "max_abs_residual_trace_diff": float(np.abs(np.array([t[3] for t in mover.trace[:m]]) - g["residual"][:m]).max())}
    cg = np.array(mover.cg_iterations)
    if rank == 0:
        rec = {"workload": f"MongeAmpereMover relaxation to rtol {a.rtol:g}, Unit{'Cube' if a.dim == 3 else 'Square'}Mesh({n}), "
                           f"{'demos/monge_ampere_3d.py' if a.dim == 3 else 'demos/monge_ampere_helmholtz.py'} monitor, through move()",
               "n_gpus": world, "vertices": int(V), "iterations_returned": int(its), "relaxation_iterations": int(its) + 1,
               "final_residual": float(mover.trace[-1][3]), "initial_residual": float(mover.trace[0][3]),
               "final_volume_ratio": float(mover.trace[-1][1]), "final_sigma_over_mu": float(mover.trace[-1][2]),
               "tangled_cells_at_convergence": int(tangled), "move_wall_s": wall, "mesh_generation_s": t_mesh, "setup_s": t_setup,
               "converged_vertex_iters_per_s": V * (int(its) + 1) / wall,
               "inner_iterations_total(l2,poisson,hessian)": [int(v) for v in cg.sum(axis=0)],
               "inner_iterations_mean(l2,poisson,hessian)": [float(v) for v in cg.mean(axis=0)],
               "x_sum": [float(v) for v in sums[: a.dim]], "x_sumsq": [float(v) for v in sums[a.dim:]],
               "compare": cmp_, "residual_trace_every_10": [float(t[3]) for t in mover.trace[::10]]}
        print(json.dumps(rec), flush=True)
        if a.out:
            with open(a.out, "w") as f:
                json.dump(rec, f, indent=1)
    parallel.finalize()


if __name__ == "__main__":
    main()
