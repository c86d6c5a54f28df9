#!/usr/bin/env python
"""bench.py -- the headline metric of BASELINE.json: Monge-Ampere relaxation vertex-iterations/s
on the 3D UnitCubeMesh (demos/monge_ampere_3d.py monitor), 1/2/4/8 B200, with the HBM-roofline
fraction of the dominant kernel and the CPU baseline timed beside it.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--n 256] [--impl reference]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is ONE relaxation iteration (movement/monge_ampere.py:500-538) of the whole mesh: L2
projection of grad(phi) (d mass solves), x = xi + grad(phi) fused with the monitor, tangling
check, theta/residual/volume diagnostics, the Poisson pseudo-timestep (PCG to rtol 1e-12 -- every
solve is converged, nothing is skipped) and the Hessian recovery (d*d mass solves).
`value` = V * K / t with the mesh resident in HBM; `e2e` = the same through the public API from
host arrays (H2D of the mesh + ingest/assembly + K iterations + D2H of the moved coordinates).
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "MA relax vertex-iters/s on 3D cube mesh"
UNIT = "vertex-iters/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--n", type=int, default=int(os.environ.get("MV_BENCH_N", "256")),
                    help="UnitCubeMesh(n,n,n); 256 is BASELINE.json configs[4] (~100M tetrahedra)")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--inner-rtol", type=float, default=1e-12)
    ap.add_argument("--cpu-n", type=int, default=int(os.environ.get("MV_BENCH_CPU_N", "48")),
                    help="cube size of the bounded CPU-baseline sample")
    ap.add_argument("--cpu-steps", type=int, default=2)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--mass-solver", default="chebyshev", choices=["chebyshev", "pcg"],
                    help="solver of the consistent-mass systems (A1, A8): fused Chebyshev (default) or Jacobi-PCG")
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------
# clocks: sampled DURING the timed region (B200_PROFILING.md)
# ---------------------------------------------------------------------------------------------
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(self.index), "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 6:
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def ncu_traffic():
    """Per-launch DRAM bytes of the dominant kernel from the committed ncu capture, if any."""
    p = os.path.join(ROOT, "profiles", "dominant_kernel.json")
    if os.path.exists(p):
        with open(p) as f:
            return json.load(f)
    return None


# ---------------------------------------------------------------------------------------------
# CPU arm: the oracle's C/OpenMP restatement (the reference's Firedrake/PETSc path is unavailable
# on this box -- BASELINE.md section 2), all host threads, a bounded sample of the same workload.
# ---------------------------------------------------------------------------------------------
def cpu_relax(n, steps, warmup, inner_rtol):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle_c as oc

    from movement_b200.mesh import cube_arrays

    coords, cells, bf, bt = cube_arrays(n, n, n)
    cores = os.cpu_count() or 1
    c = oc.CRelaxation(coords, cells, bf, bt, oc.MON_SINATAN3, (10.0,), inner_rtol=inner_rtol, nthreads=cores)
    for _ in range(warmup):
        c.step(update=True)
    t0 = time.perf_counter()
    its = []
    for _ in range(steps):
        _, it, _ = c.step(update=True)
        its.append([int(v) for v in it])
    dt = time.perf_counter() - t0
    V = coords.shape[0]
    return {"value": V * steps / dt, "unit": UNIT, "cores": int(c.threads), "kind": "port",
            "sample": f"UnitCubeMesh({n}^3): V={V}, C={cells.shape[0]}, {steps} relaxation iterations after {warmup} warm-up, "
                      f"inner Jacobi-PCG rtol {inner_rtol:g}, C/OpenMP oracle port (oracle/movement_oracle_c.c)",
            "seconds": dt, "cg_iterations": its}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # bounded: the CPU port needs ~9n Poisson CG iterations per step; n=48 keeps K+W steps within minutes
    r = cpu_relax(args.cpu_n, max(args.steps, 1), min(args.warmup, 1), args.inner_rtol)
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": min(args.warmup, 1), "ms_per_step": 1e3 * r["seconds"] / max(args.steps, 1), "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(args.n), "n": args.n, "sample_n": args.cpu_n},
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "reference's own Firedrake/PETSc path unavailable on this box; this is the oracle's C/OpenMP port"}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------
def workload_name(n):
    return (f"MongeAmpereMover relaxation (pseudo_timestep 0.1, inner solves rtol 1e-12), demos/monge_ampere_3d.py monitor, "
            f"UnitCubeMesh({n},{n},{n}) = BASELINE.json configs[4]" if n == 256 else
            f"MongeAmpereMover relaxation, demos/monge_ampere_3d.py monitor, UnitCubeMesh({n},{n},{n}) (reduced size of configs[4])")


def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    import movement_b200 as mv
    from movement_b200 import _lib as L
    from movement_b200 import parallel
    from movement_b200.mesh import Mesh, cube_arrays

    assert torch.cuda.is_available(), "bench.py needs a CUDA device: movement_b200 has no CPU fallback"
    parallel.init()
    rank, world = parallel.rank(), parallel.size()
    dev = torch.cuda.current_device()
    lib = L.load()
    cheb = args.mass_solver == "chebyshev"
    L.check(lib.mv_set_mass_solver(L.MV_MASS_CHEBYSHEV if cheb else L.MV_MASS_PCG))
    n = args.n
    quiet = mv.mover  # silence the per-iteration prints inside the timed region
    quiet._print = lambda *a, **k: None
    mv.monge_ampere._print = quiet._print
    mv.monge_ampere.MongeAmpereMover_Base.inner_rtol = args.inner_rtol

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- host mesh (the caller's input): pinned host arrays ------------------------------------
    arrays = cube_arrays(n, n, n)
    if world > 1:
        from movement_b200.partition import partition_mesh

        lp = partition_mesh(*arrays, rank, world)
        host = (lp.coords, lp.cells, lp.bfacets, lp.btags)
        Vglobal, Cglobal = lp.global_num_vertices, lp.global_num_cells
    else:
        lp = None
        host = arrays
        Vglobal, Cglobal = arrays[0].shape[0], arrays[1].shape[0]
    del arrays
    # the caller's buffers live in pinned host memory (numpy views of pinned torch tensors)
    pinned = [torch.from_numpy(np.ascontiguousarray(a)).pin_memory() for a in host]
    host = tuple(t.numpy() for t in pinned)
    h2d_bytes = sum(int(a.nbytes) for a in host)

    def make_mover():
        mesh = Mesh(*host, part=lp)          # H2D of the coordinates; topology copied at ingest
        return mv.MongeAmpereMover(mesh, mv.sinatan3_hessian_monitor(), method="relaxation", rtol=1e-8,
                                   maxiter=10 ** 6, dtol=1e300)

    def step(mover):
        r = mover._evaluate()
        mover._advance_potential()
        return r

    # ---- resident-in-HBM measurement -------------------------------------------------------------
    mover = make_mover()
    mover._begin()
    for _ in range(args.warmup):
        step(mover)
    L.check(lib.mv_prof_reset())
    L.check(lib.mv_prof_enable(1))
    sampler = ClockSampler(dev)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    residuals = [step(mover) for _ in range(args.steps)]
    e1.record()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    L.check(lib.mv_prof_enable(0))
    spmv_ms = (C.c_double * 16)()
    spmv_cnt = (C.c_int64 * 16)()
    pcg_it = (C.c_int64 * 16)()
    launches = C.c_int64(0)
    L.check(lib.mv_prof_get(spmv_ms, spmv_cnt, pcg_it, C.byref(launches)))
    cg = np.array(mover.cg_iterations[-args.steps:])
    value = Vglobal * args.steps / (ms * 1e-3)

    # ---- roofline of the dominant kernel: the matrix-streaming kernels of the inner solves, timed live per launch with
    # CUDA events on the launching stream (mv_prof_*), by number of right-hand sides C: 3 = L2 projection (mass),
    # 1 = Poisson (stiffness, SpMV of the multigrid-preconditioned CG), 9 = Hessian recovery (mass).
    # Algorithmic bytes per launch: 12 B per stored non-zero + 4 B slice pointer per row, plus
    #   SpMV (+ fused p.q):            read p, write q                               = 16 C B per row
    #   fused Chebyshev iteration:     dinv, b, x_{k-1} in, x_{k+1} out, own x_k     = 8 + 32 C B per row (+2 B mask)
    dm = mover._dm
    peak, peak_src = peaks()
    have_mask = bool(getattr(mover, "_have_bc", True))
    classes = {}
    for c in (1, 3, 9):
        if spmv_cnt[c] > 0:
            fused = cheb and c != 1
            per_row = (4.0 + 8.0 + 32.0 * c + (2.0 if (c == 3 and have_mask) else 0.0)) if fused else (4.0 + 16.0 * c)
            alg = 12.0 * dm.nnz + per_row * dm.Vown
            avg = spmv_ms[c] / spmv_cnt[c]
            classes[c] = {"rhs": c, "kernel": ("k_cheb_tma" if c == 9 else "k_cheb_soa") if fused else "k_spmv_soa", "launches_timed": int(spmv_cnt[c]),
                          "avg_launch_ms": avg, "algorithmic_bytes_per_launch": alg,
                          "achieved_GBs": alg / (avg * 1e-3) / 1e9, "share_of_step": spmv_ms[c] / ms}
    dom = max(classes, key=lambda c: classes[c]["share_of_step"])
    traffic = ncu_traffic()
    if cheb:
        kname = {1: "SELL-32 SpMV + fused p.q, k_spmv_soa<1,true> (stiffness matrix, Poisson PCG)",
                 3: "fused Chebyshev iteration, k_cheb_soa<3,...> (mass matrix, L2 projection of grad phi)",
                 9: "fused Chebyshev iteration, TMA-staged matrix stream, k_cheb_tma<9,3,4,3,2> (mass matrix, Hessian recovery)"}[dom]
    else:
        kname = "SELL-32 SpMV + fused p.q, " + {1: "k_spmv_soa<1,true> (stiffness matrix, Poisson PCG)",
                                                 3: "k_spmv_soa<3,true> (mass matrix, L2 projection of grad phi)",
                                                 9: "k_spmv_soa<9,true> (mass matrix, Hessian recovery)"}[dom]
    tr = (traffic or {}).get(("cheb" if cheb and dom != 1 else "spmv") + str(dom)) or {}
    roofline = {"bound": "hbm", "kernel": kname,
                "achieved": classes[dom]["achieved_GBs"], "peak": peak, "unit": "GB/s", "frac": classes[dom]["achieved_GBs"] / peak,
                "peak_source": peak_src, "frac_of_nominal_8TBs": classes[dom]["achieved_GBs"] / 8000.0,
                "algorithmic_bytes_per_launch": classes[dom]["algorithmic_bytes_per_launch"],
                "avg_launch_ms": classes[dom]["avg_launch_ms"], "launches_timed": classes[dom]["launches_timed"],
                "share_of_step": classes[dom]["share_of_step"],
                "traffic": tr.get("dram_bytes_per_launch"), "traffic_source": tr.get("source"),
                "all_matrix_streaming_classes": [classes[c] for c in sorted(classes)]}
    # whole-step accounting with the SURVEY 8d byte model per vertex per inner iteration: B_cg(c) = 12 nnz + 4 + 104 c
    # for a CG iteration, B_cheb(c) = 12 nnz + 12 + 32 c for a fused Chebyshev iteration
    nnzv = dm.nnz / dm.Vown
    d = 3
    bcg = lambda c: 12 * nnzv + 4 + 104 * c  # noqa: E731
    bmass = (lambda c: 12 * nnzv + 12 + 32 * c) if cheb else bcg  # noqa: E731
    step_bytes = float(np.mean(cg[:, 0] * bmass(d) + cg[:, 1] * bcg(1) + cg[:, 2] * bmass(d * d))) * Vglobal
    step_gbs = step_bytes / (ms / args.steps * 1e-3) / 1e9

    # ---- where a step goes: two more (untimed above) iterations with CUDA events between the phases, max over ranks
    def phase_step(mv_):
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
        ev[0].record()
        n1 = mv_._l2_project()
        ev[1].record()
        mv_._update_physical_coordinates()
        mv_._n1 = n1
        mv_._diagnostics(use_H_old=True)
        ev[2].record()
        n2 = mv_._pseudotimestep()
        ev[3].record()
        n3 = mv_._equidistribute()
        mv_._advance()
        mv_.cg_iterations.append((n1, n2, n3))
        ev[4].record()
        torch.cuda.synchronize()
        return [ev[i].elapsed_time(ev[i + 1]) for i in range(4)]

    ph = torch.tensor([phase_step(mover) for _ in range(2)], dtype=torch.float64, device="cuda").mean(dim=0)
    if world > 1:
        dist.all_reduce(ph, op=dist.ReduceOp.MAX)
    phases = dict(zip(["l2_projection(A1)", "update_x+monitor+tangling+diagnostics(A2-A6)", "poisson_pseudotimestep(A7)",
                       "hessian_recovery(A8)"], [float(v) for v in ph.tolist()]))

    # ---- end to end through the public API from host buffers ---------------------------------------
    e2e = None
    if not args.no_e2e:
        del mover
        torch.cuda.empty_cache()
        out_host = torch.empty((host[0].shape[0], 3), dtype=torch.float64).pin_memory()   # the caller's result buffer
        barrier()
        t0 = time.perf_counter()
        m2 = make_mover()
        torch.cuda.synchronize(); t1 = time.perf_counter()
        m2._begin()
        for _ in range(args.steps):
            step(m2)
        torch.cuda.synchronize(); t2 = time.perf_counter()
        m2.to_physical_coordinates()
        out = out_host[: m2._dm.Vown]
        out.copy_(m2.mesh.coordinates.tensor[: m2._dm.Vown], non_blocking=True)   # D2H of the moved (owned) coordinates
        torch.cuda.synchronize()
        barrier()
        t3 = time.perf_counter()
        t = torch.tensor([t3 - t0], dtype=torch.float64, device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e = {"value": Vglobal * args.steps / float(t.item()), "unit": UNIT,
               "h2d_bytes_per_step": h2d_bytes / args.steps, "d2h_bytes_per_step": (out.numel() * 8 + 6 * 8 * args.steps) / args.steps,
               "seconds": float(t.item()),
               "breakdown_s": {"h2d+ingest+assembly": t1 - t0, "iterations (incl. multigrid set-up)": t2 - t1, "d2h": t3 - t2},
               "what": "MongeAmpereMover(host mesh).move-equivalent: H2D mesh + ingest/assembly + K iterations + D2H coordinates"}
        del m2

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic",
                "config": {"workload": workload_name(n), "n": n, "vertices": Vglobal, "cells": Cglobal,
                           "parallelism": f"vertex-owner slab partition x{world}" if world > 1 else "single GPU",
                           "l2": "inputs larger than L2 (matrix stream %.1f GB per SpMV)" % (12.0 * dm.nnz / 1e9),
                           "mass_solver": "fused Chebyshev (Wathen interval), measured stop at rtol 1e-12" if cheb else "Jacobi-PCG",
                           "inner_iterations_per_step(l2,poisson,hessian)": [float(x) for x in cg.mean(axis=0)],
                           "residual_after_timed_steps": residuals[-1]},
                "roofline": roofline,
                "step_model": {"bytes_per_step": step_bytes, "GBs": step_gbs, "frac_of_peak": step_gbs / (peak * world)},
                "phases_ms": phases,
                "gpu_launches": int(launches.value), "clocks": clocks, "e2e": e2e}
        if world == 1 and not args.no_cpu_baseline:
            r = cpu_relax(args.cpu_n, args.cpu_steps, 1, args.inner_rtol)
            line["cpu_baseline"] = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")}
        print(json.dumps(line), flush=True)
    parallel.finalize()


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
