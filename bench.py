#!/usr/bin/env python
"""bench.py -- the headline metric of BASELINE.json: Monge-Ampere relaxation vertex-iterations/s on the 3D
UnitCubeMesh (demos/monge_ampere_3d.py monitor), 1/2/4/8 B200, with the HBM-roofline fraction of the dominant
kernel and the CPU baseline timed beside it.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--n 256] [--config 5|2|3|4] [--impl reference]
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

config 5 (default, the headline; BASELINE.json configs[4]).  A "step" is ONE relaxation iteration
(movement/monge_ampere.py:500-538) of the whole mesh: L2 projection of grad(phi) (d mass solves), x = xi + grad(phi)
fused with the monitor, theta/residual/volume diagnostics, the Poisson pseudo-timestep and the Hessian recovery
(d*d mass solves), every inner solve converged to rtol 1e-12 (nothing skipped).
  value = V * K / t, K iterations of the device-side loop (mv_ma_relax_move, the C-ABI call behind move()) with the
          mesh resident in HBM;
  e2e   = the same through the public API from pinned host arrays: Mesh(host arrays) -> MongeAmpereMover(...) ->
          .move() (maxiter = K + 1: K full iterations and the evaluation half of the next, as the reference's loop
          does before it reports "failed to converge") -> D2H of the moved coordinates; counted as K.
  parity  = residual / theta trace and sampled coordinates of the first 5 iterations of a fresh mover against the
          stored single-GPU record (tests/golden/bench_parity_n<n>.npz, written with --write-parity at N = 1), so a
          multi-GPU line carries a parity number and not only a speed.
  true_relres_max = Euclidean ||b - K x|| / ||b|| of the iterates the Poisson PCG returned, recomputed at exit.

configs 2, 3, 4 print one line each for the other BASELINE configs (Laplacian smoothing + tangling 1024^2, lineal
springs 4096^2, 2D Helmholtz relaxation / quasi-Newton 8192^2) with SURVEY 8d's metrics for them.
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
import warnings

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "MA relax vertex-iters/s on 3D cube mesh"
UNIT = "vertex-iters/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--config", type=int, default=5, choices=[2, 3, 4, 5], help="BASELINE.json configs[config-1]")
    ap.add_argument("--n", type=int, default=int(os.environ.get("MV_BENCH_N", "0")),
                    help="mesh size per direction (default: the BASELINE size of the config: 256 / 1024 / 4096 / 8192)")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--method", default="relaxation", choices=["relaxation", "quasi_newton"], help="config 4 only")
    ap.add_argument("--inner-rtol", type=float, default=1e-12)
    ap.add_argument("--cpu-n", type=int, default=int(os.environ.get("MV_BENCH_CPU_N", "0")),
                    help="cube size of the bounded CPU sample (default 96 inside the GPU line, 128 for --impl reference)")
    ap.add_argument("--cpu-steps", type=int, default=2)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--e2e-detail", action="store_true", help="one library call per iteration in the e2e leg, wall time of each reported")
    ap.add_argument("--write-parity", action="store_true", help="N = 1: store the parity record of this n")
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
def cpu_relax(n, steps, warmup, inner_rtol, dim=3):
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle_c as oc

    from movement_b200.mesh import cube_arrays, square_arrays

    coords, cells, bf, bt = cube_arrays(n, n, n) if dim == 3 else square_arrays(n, n)
    cores = os.cpu_count() or 1
    kind, par = (oc.MON_SINATAN3, (10.0,)) if dim == 3 else (oc.MON_GAUSS_HESS, (0.51, 0.65, 0.1, 5.0))
    c = oc.CRelaxation(coords, cells, bf, bt, kind, par, inner_rtol=inner_rtol, nthreads=cores)
    for _ in range(warmup):
        c.step(update=True)
    t0 = time.perf_counter()
    its = []
    for _ in range(steps):
        _, it, _ = c.step(update=True)
        its.append([int(v) for v in it])
    dt = time.perf_counter() - t0
    V = coords.shape[0]
    mesh = f"UnitCubeMesh({n}^3)" if dim == 3 else f"UnitSquareMesh({n}^2)"
    return {"value": V * steps / dt, "unit": UNIT, "cores": int(c.threads), "kind": "port",
            "sample": f"{mesh}: V={V}, C={cells.shape[0]}, {steps} relaxation iterations after {warmup} warm-up, "
                      f"inner Jacobi-PCG rtol {inner_rtol:g}, C/OpenMP oracle port (oracle/movement_oracle_c.c)",
            "n": n, "seconds": dt, "cg_iterations": its}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    # bounded: a step of the CPU port costs ~9n Jacobi-PCG Poisson iterations; 128^3 keeps K + W steps within minutes
    if args.config == 4:
        n, cpu_n, dim = args.n or 8192, args.cpu_n or 512, 2
    else:
        n, cpu_n, dim = args.n or 256, args.cpu_n or 128, 3
    wu = min(args.warmup, 1)
    # never exit without a line: if the requested sample cannot be set up or run on this host (memory, an assertion in
    # the port), halve it and say so in the line
    r, fallback = None, ""
    while r is None:
        try:
            r = cpu_relax(cpu_n, max(args.steps, 1), wu, args.inner_rtol, dim=dim)
        except Exception as exc:  # noqa: BLE001
            if cpu_n <= 8:
                raise
            fallback += f"sample {cpu_n}^{dim} failed ({type(exc).__name__}: {str(exc)[:120]}), retried at {cpu_n // 2}^{dim}; "
            cpu_n //= 2
    line = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": wu, "ms_per_step": 1e3 * r["seconds"] / max(args.steps, 1), "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": workload_name(n, dim), "n": n, "sample_n": cpu_n},
            "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample")},
            "e2e": {"value": r["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": fallback + "the reference's own Firedrake/PETSc path is unavailable on this box; this is the oracle's C/OpenMP port "
                    f"(Jacobi-PCG inner solves) on a bounded {cpu_n}^{dim} sample of the workload -- a different size and a weaker "
                    "Poisson solver than the GPU arm (multigrid PCG); the GPU line carries the same-size, same-algorithm pair "
                    "in cpu_baseline.gpu_same_sample"}
    print(json.dumps(line), flush=True)


# ---------------------------------------------------------------------------------------------
# GPU arm, common pieces
# ---------------------------------------------------------------------------------------------
def workload_name(n, dim=3):
    if dim == 3:
        return (f"MongeAmpereMover relaxation (pseudo_timestep 0.1, inner solves rtol 1e-12), demos/monge_ampere_3d.py monitor, "
                f"UnitCubeMesh({n},{n},{n}) = BASELINE.json configs[4]" if n == 256 else
                f"MongeAmpereMover relaxation, demos/monge_ampere_3d.py monitor, UnitCubeMesh({n},{n},{n}) (reduced size of configs[4])")
    return (f"MongeAmpereMover, demos/monge_ampere_helmholtz.py monitor, UnitSquareMesh({n},{n})"
            + (" = BASELINE.json configs[3]" if n == 8192 else " (reduced size of configs[3])"))


class Ctx:
    """torch / library handles shared by the config runners."""

    def __init__(self, args):
        import numpy as np
        import torch
        import torch.distributed as dist

        import movement_b200 as mv
        from movement_b200 import _lib as L
        from movement_b200 import parallel

        assert torch.cuda.is_available(), "bench.py needs a CUDA device: movement_b200 has no CPU fallback"
        parallel.init()
        self.np, self.torch, self.dist, self.mv, self.L, self.parallel = np, torch, dist, mv, L, parallel
        self.rank, self.world = parallel.rank(), parallel.size()
        self.dev = torch.cuda.current_device()
        self.lib = L.load()
        self.args = args
        quiet = lambda *a, **k: None  # noqa: E731  (no per-iteration prints inside timed regions)
        mv.mover._print = quiet
        mv.monge_ampere._print = quiet
        mv.laplacian_smoothing._print = quiet
        mv.spring._print = quiet
        warnings.simplefilter("ignore")

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, v):
        t = self.torch.tensor([float(v)], dtype=self.torch.float64, device="cuda")
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def pinned(self, arrays):
        ts = [self.torch.from_numpy(self.np.ascontiguousarray(a)).pin_memory() for a in arrays]
        return tuple(t.numpy() for t in ts), ts

    def prof(self):
        spmv_ms, spmv_cnt, pcg_it = (C.c_double * 16)(), (C.c_int64 * 16)(), (C.c_int64 * 16)()
        launches = C.c_int64(0)
        self.L.check(self.lib.mv_prof_get(spmv_ms, spmv_cnt, pcg_it, C.byref(launches)))
        return list(spmv_ms), list(spmv_cnt), list(pcg_it), int(launches.value)

    def true_res(self):
        mx, last, cnt = (C.c_double * 16)(), (C.c_double * 16)(), (C.c_int64 * 16)()
        self.L.check(self.lib.mv_true_residual_get(mx, last, cnt))
        return list(mx), list(cnt)


def host_mesh(cx, arrays_fn, part_fn=None):
    """The caller's input: (pinned) host arrays of this rank's part, with the time it took to make them.  On several GPUs
    the generated meshes are created slab by slab (mesh.cube_part / square_part: identical to cutting the global mesh,
    tests/test_partition.py) unless MV_PARTITION selects another partitioner."""
    t0 = time.perf_counter()
    lp = None
    if cx.world > 1 and part_fn is not None and os.environ.get("MV_PARTITION", "block") == "block":
        lp = part_fn(cx.rank, cx.world)
        t1 = t2 = time.perf_counter()
        arrays = (lp.coords, lp.cells, lp.bfacets, lp.btags)
        Vg, Cg = lp.global_num_vertices, lp.global_num_cells
    else:
        arrays = arrays_fn()
        t1 = time.perf_counter()
        Vg, Cg = arrays[0].shape[0], arrays[1].shape[0]
        if cx.world > 1:
            from movement_b200.partition import partition_mesh

            lp = partition_mesh(*arrays, cx.rank, cx.world)
            arrays = (lp.coords, lp.cells, lp.bfacets, lp.btags)
        t2 = time.perf_counter()
    host, keep = cx.pinned(arrays)
    return host, keep, lp, Vg, Cg, {"mesh_generation_s": t1 - t0, "partition_s": t2 - t1, "pin_s": time.perf_counter() - t2}


def relax_k(cx, mover, k, i0):
    """k iterations of the device-side loop (the C-ABI call behind move()), no exit tests firing."""
    L = cx.L
    trace, its, info = (C.c_double * (6 * k))(), (C.c_int * (3 * k))(), (C.c_int * 3)()
    res0 = C.c_double(1.0)
    L.check(cx.lib.mv_ma_relax_move(mover._ma, mover.pseudo_dt, i0, k, 2 ** 30, 0.0, 1e300, C.byref(res0), trace, its, info))
    tr = cx.np.array(trace[:]).reshape(k, 6)
    return tr, cx.np.array(its[:]).reshape(k, 3)


def move_k(mover, k):
    """mover.move() for k full iterations (maxiter = k + 1; the reference's loop raises after the evaluation half of
    iteration k + 1)."""
    mover.maxiter = k + 1
    mover.rtol = 0.0
    mover.dtol = 1e300
    try:
        mover.move()
    except Warning:
        pass
    return mover


# ---------------------------------------------------------------------------------------------
# config 5 (headline) and config 4 relaxation: Monge-Ampere relaxation
# ---------------------------------------------------------------------------------------------
def run_relaxation(args, dim):
    cx = Ctx(args)
    np, torch, mv, L, lib = cx.np, cx.torch, cx.mv, cx.L, cx.lib
    from movement_b200.mesh import Mesh, cube_arrays, cube_part, square_arrays, square_part

    rank, world = cx.rank, cx.world
    cheb = args.mass_solver == "chebyshev"
    L.check(lib.mv_set_mass_solver(L.MV_MASS_CHEBYSHEV if cheb else L.MV_MASS_PCG))
    n = args.n or (256 if dim == 3 else 8192)
    mv.monge_ampere.MongeAmpereMover_Base.inner_rtol = args.inner_rtol
    monitor = mv.sinatan3_hessian_monitor() if dim == 3 else mv.helmholtz_hessian_monitor()
    host, keep, lp, Vglobal, Cglobal, prep = host_mesh(cx, (lambda: cube_arrays(n, n, n)) if dim == 3 else (lambda: square_arrays(n, n)),
                                                       (lambda r, s: cube_part(n, n, n, r, s)) if dim == 3 else (lambda r, s: square_part(n, n, r, s)))
    h2d_bytes = sum(int(a.nbytes) for a in host)

    def make_mover(maxiter=10 ** 6):
        mesh = Mesh(*host, part=lp)          # H2D of the coordinates; topology copied at ingest
        return mv.MongeAmpereMover(mesh, monitor, method="relaxation", rtol=0.0, maxiter=maxiter, dtol=1e300,
                                   raise_convergence_errors=False)

    # ---- resident-in-HBM measurement -------------------------------------------------------------
    mover = make_mover()
    mover._begin()
    if args.warmup > 0:
        relax_k(cx, mover, args.warmup, 0)
    L.check(lib.mv_prof_reset())
    L.check(lib.mv_prof_enable(1))
    sampler = ClockSampler(cx.dev)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    cx.barrier()
    e0.record()
    trace, cg = relax_k(cx, mover, args.steps, args.warmup)
    e1.record()
    cx.barrier()
    clocks = sampler.stop() if rank == 0 else None
    ms = cx.max_over_ranks(e0.elapsed_time(e1))
    L.check(lib.mv_prof_enable(0))
    spmv_ms, spmv_cnt, pcg_it, launches = cx.prof()
    value = Vglobal * args.steps / (ms * 1e-3)

    # ---- roofline of the dominant kernel: the matrix-streaming kernels of the inner solves, timed live per launch with
    # CUDA events on the launching stream (mv_prof_*), by number of right-hand sides C: d = L2 projection (mass),
    # 1 = Poisson (stiffness, SpMV of the multigrid-preconditioned CG), d*d = Hessian recovery (mass).
    # Algorithmic bytes per launch: 12 B per stored non-zero + 4 B slice pointer per row, plus
    #   SpMV (+ fused p.q):            read p, write q                               = 16 C B per row
    #   fused Chebyshev iteration:     dinv, b, x_{k-1} in, x_{k+1} out, own x_k     = 8 + 32 C B per row (+2 B mask)
    dm = mover._dm
    peak, peak_src = peaks()
    have_mask = bool(getattr(mover, "_have_bc", True))
    classes = {}
    mixed = cheb and os.environ.get("MV_CHEB_MIXED", "1") != "0"
    #   mixed-precision Chebyshev (default): single-precision iteration: 8 B per non-zero; dinv, r, d_{k-1} in, d_{k+1} out,
    #   own d_k = 4 + 16 C B per row; double-precision residual pass (timed as class C + 4): 12 B per non-zero; dinv, b, x in,
    #   r and d_1 out (single) = 8 + 24 C B per row
    for c in (1, dim, dim * dim):
        if spmv_cnt[c] > 0:
            fused = cheb and c != 1
            if fused and mixed:
                alg = 8.0 * dm.nnz + (4.0 + 4.0 + 16.0 * c + (2.0 if (c == dim and have_mask) else 0.0)) * dm.Vown
                kern = ("k_cheb_tma<...,float,float,MIX_IT4> (fp32 Chebyshev iteration, TMA-staged stream, packed float4 iterates)" if c == 9
                        else "k_mcheb (fp32 Chebyshev iteration)")
            else:
                per_row = (4.0 + 8.0 + 32.0 * c + (2.0 if (c == dim and have_mask) else 0.0)) if fused else (4.0 + 16.0 * c)
                alg = 12.0 * dm.nnz + per_row * dm.Vown
                kern = ("k_cheb_tma" if c == 9 else "k_cheb_soa") if fused else "k_spmv_soa"
            avg = spmv_ms[c] / spmv_cnt[c]
            classes[c] = {"rhs": c, "kernel": kern, "launches_timed": int(spmv_cnt[c]),
                          "avg_launch_ms": avg, "algorithmic_bytes_per_launch": alg,
                          "achieved_GBs": alg / (avg * 1e-3) / 1e9, "share_of_step": spmv_ms[c] / ms}
            if fused and mixed and spmv_cnt[c + 4] > 0:
                alg = 12.0 * dm.nnz + (4.0 + 8.0 + 24.0 * c + (2.0 if (c == dim and have_mask) else 0.0)) * dm.Vown
                avg = spmv_ms[c + 4] / spmv_cnt[c + 4]
                classes[c + 4] = {"rhs": c, "kernel": "fp64 residual pass of the mixed-precision Chebyshev solver", "launches_timed": int(spmv_cnt[c + 4]),
                                  "avg_launch_ms": avg, "algorithmic_bytes_per_launch": alg,
                                  "achieved_GBs": alg / (avg * 1e-3) / 1e9, "share_of_step": spmv_ms[c + 4] / ms}
    dom = max(classes, key=lambda c: classes[c]["share_of_step"])
    traffic = ncu_traffic()
    domc = classes[dom]["rhs"]
    role = {1: "stiffness matrix, Poisson PCG", dim: "mass matrix, L2 projection of grad phi", dim * dim: "mass matrix, Hessian recovery"}[domc]
    if cheb and domc != 1 and mixed:
        kname = f"{classes[dom]['kernel']}, C = {domc} ({role})"
    elif cheb and domc != 1:
        kname = f"fused Chebyshev iteration, {'TMA-staged matrix stream, k_cheb_tma' if dom == 9 else 'k_cheb_soa'}<{dom},...> ({role})"
    else:
        kname = f"SELL-32 SpMV + fused p.q, k_spmv_soa<{dom},true> ({role})"
    tr = (traffic or {}).get(("mcheb" if mixed and domc != 1 else ("cheb" if cheb and domc != 1 else "spmv")) + str(domc)) or {}
    roofline = {"bound": "hbm", "kernel": kname,
                "achieved": classes[dom]["achieved_GBs"], "peak": peak, "unit": "GB/s", "frac": classes[dom]["achieved_GBs"] / peak,
                "peak_source": peak_src, "frac_of_nominal_8TBs": classes[dom]["achieved_GBs"] / 8000.0,
                "algorithmic_bytes_per_launch": classes[dom]["algorithmic_bytes_per_launch"],
                "avg_launch_ms": classes[dom]["avg_launch_ms"], "launches_timed": classes[dom]["launches_timed"],
                "share_of_step": classes[dom]["share_of_step"],
                "traffic": tr.get("dram_bytes_per_launch") if (dim == 3 and n == 256 and world == 1) else None,
                "traffic_source": tr.get("source"),
                "all_matrix_streaming_classes": [classes[c] for c in sorted(classes)]}
    # whole-step accounting with the SURVEY 8d byte model per vertex per inner iteration: B_cg(c) = 12 nnz + 4 + 104 c
    # for a CG iteration, B_cheb(c) = 12 nnz + 12 + 32 c for a fused Chebyshev iteration
    nnzv = dm.nnz / dm.Vown
    d = dim
    bcg = lambda c: 12 * nnzv + 4 + 104 * c  # noqa: E731
    bmass = (lambda c: 12 * nnzv + 12 + 32 * c) if cheb else bcg  # noqa: E731
    step_bytes = float(np.mean(cg[:, 0] * bmass(d) + cg[:, 1] * bcg(1) + cg[:, 2] * bmass(d * d))) * Vglobal
    if mixed:   # mass solves: the launches of the timed region with their own byte counts (this rank's share x world)
        step_bytes = float(np.mean(cg[:, 1] * bcg(1))) * Vglobal + world * sum(
            v["algorithmic_bytes_per_launch"] * v["launches_timed"] for k, v in classes.items() if v["rhs"] != 1) / args.steps
    step_gbs = step_bytes / (ms / args.steps * 1e-3) / 1e9

    # ---- where a step goes: two more (untimed above) iterations with CUDA events between the phases, max over ranks;
    # the Poisson solves of these two are audited (true residual of the returned iterate)
    L.check(lib.mv_true_residual_enable(1))

    def phase_step(mv_):
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
        ev[0].record()
        n1 = mv_._l2_project()
        ev[1].record()
        mv_._sync()
        mv_._call(lib.mv_ma_update_coordinates)
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
        cx.dist.all_reduce(ph, op=cx.dist.ReduceOp.MAX)
    phases = dict(zip(["l2_projection(A1)", "update_x+monitor+diagnostics(A2-A6)", "poisson_pseudotimestep(A7)",
                       "hessian_recovery(A8)"], [float(v) for v in ph.tolist()]))
    tmax, tcnt = cx.true_res()
    L.check(lib.mv_true_residual_enable(0))
    true_relres = {"poisson_pcg_max": tmax[1], "solves_audited": int(tcnt[1]), "rtol": args.inner_rtol,
                   "what": "Euclidean ||b - K x|| / ||b|| of the returned iterate, one extra SpMV at PCG exit (mv_true_residual_*)"}
    amg = dm.amg_info()
    l2_gb = 12.0 * dm.nnz / 1e9
    # the resident mover AND its device mesh go away before the end-to-end leg: the stream-ordered pool keeps their pages,
    # so the next mover of the process starts warm (the designed steady state of a process that moves mesh after mesh)
    del mover, dm
    import gc

    gc.collect()
    torch.cuda.empty_cache()

    # ---- end to end through the public API from host buffers: Mesh(host) -> MongeAmpereMover -> move() -> D2H ----------
    e2e = None
    if not args.no_e2e:
        out_host = torch.empty((host[0].shape[0], dim), dtype=torch.float64).pin_memory()   # the caller's result buffer
        cx.barrier()
        t0 = time.perf_counter()
        m2 = make_mover()
        torch.cuda.synchronize(); t1 = time.perf_counter()
        if args.e2e_detail:
            m2.move_chunk = 1
        move_k(m2, args.steps)
        torch.cuda.synchronize(); t2 = time.perf_counter()
        out = out_host[: m2._dm.Vown]
        out.copy_(m2.x.tensor[: m2._dm.Vown], non_blocking=True)   # D2H of the moved (owned) coordinates
        torch.cuda.synchronize()
        cx.barrier()
        t3 = time.perf_counter()
        t = cx.max_over_ranks(t3 - t0)
        e2e = {"value": Vglobal * args.steps / t, "unit": UNIT,
               "h2d_bytes_per_step": h2d_bytes / args.steps, "d2h_bytes_per_step": (out.numel() * 8 + 6 * 8 * args.steps) / args.steps,
               "seconds": t, "iterations_counted": args.steps,
               "breakdown_s": {"Mesh(host)+MongeAmpereMover(): h2d, ingest, assembly": t1 - t0,
                               "move() (incl. multigrid set-up)": t2 - t1, "d2h": t3 - t2},
               "outside_the_timer_s": prep,
               "move_chunks(last_iteration,seconds_since_loop_start)": [(int(i), float(t)) for i, t in getattr(m2, "chunk_wall", [])],
               "what": "Mesh(pinned host arrays) -> MongeAmpereMover(mesh, monitor, method='relaxation', maxiter=K+1, "
                       "raise_convergence_errors=False) -> .move() -> D2H of mover.x; K full iterations + the evaluation half "
                       "of iteration K+1 are inside the timer and K is counted"}
        del m2
        torch.cuda.empty_cache()

    # ---- parity record: first 5 iterations of a fresh mover against the stored single-GPU run -------------------------
    parity = None
    if not args.no_parity:
        pfile = os.path.join(ROOT, "tests", "golden", f"bench_parity_{dim}d_n{n}.npz")
        m3 = move_k(make_mover(), 5)
        tr3 = np.array(m3.trace)                       # (i, volume ratio, sigma/mu, residual, theta) x 6
        gv = lp.global_vertex[: m3._dm.Vown] if lp is not None else np.arange(Vglobal)
        stride = max(1, Vglobal // 8192)
        sel = np.nonzero(gv % stride == 0)[0]
        xs = m3.x.tensor[torch.as_tensor(sel, device="cuda")].cpu().numpy()
        if args.write_parity and world == 1:
            np.savez_compressed(pfile, n=n, dim=dim, residual=tr3[:, 3], theta=tr3[:, 4], sample_stride=stride, x_sample=xs)
        if os.path.exists(pfile):
            g = np.load(pfile)
            dx = float(np.abs(xs - g["x_sample"][gv[sel] // stride]).max()) if len(sel) else 0.0
            dx = cx.max_over_ranks(dx)
            parity = {"against": f"single-GPU record tests/golden/{os.path.basename(pfile)}", "iterations": 6,
                      "max_abs_residual_diff": float(np.abs(tr3[:, 3] - g["residual"]).max()),
                      "max_abs_theta_diff": float(np.abs(tr3[:, 4] - g["theta"]).max()),
                      "max_abs_x_diff_sampled": dx, "vertices_sampled": int(len(g["x_sample"])),
                      "residual_trace": [float(v) for v in tr3[:, 3]]}
        else:
            parity = {"against": None, "residual_trace": [float(v) for v in tr3[:, 3]],
                      "note": f"no stored record for n = {n}; run `python bench.py --n {n} --write-parity` on one GPU"}
        del m3
        torch.cuda.empty_cache()

    line = None
    if rank == 0:
        line = {"metric": METRIC if dim == 3 else "MA relax vertex-iters/s on 2D square mesh (Helmholtz monitor)",
                "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f64", "data": "synthetic",
                "config": {"workload": workload_name(n, dim), "n": n, "vertices": Vglobal, "cells": Cglobal,
                           "parallelism": f"vertex-owner slab partition x{world}" if world > 1 else "single GPU",
                           "l2": "inputs larger than L2 (matrix stream %.1f GB per SpMV)" % l2_gb,
                           "mass_solver": ("fused Chebyshev (Wathen interval), " + ("mixed precision: fp64 residual passes with the measured stopping test at rtol 1e-12, fp32 sweeps on the correction" if mixed else "all fp64, measured stop at rtol 1e-12")) if cheb else "Jacobi-PCG",
                           "poisson_solver": (f"PCG (CUDA-graph iteration) + smoothed-aggregation multigrid, {amg['levels']} levels, rows {amg['rows']}, "
                                              "Galerkin initial guess from the last 3 solves") if amg else "Jacobi-PCG",
                           "inner_iterations_per_step(l2,poisson,hessian)": [float(x) for x in cg.mean(axis=0)],
                           "residual_after_timed_steps": float(trace[-1, 1])},
                "roofline": roofline,
                "step_model": {"bytes_per_step": step_bytes, "GBs": step_gbs, "frac_of_peak": step_gbs / (peak * world)},
                "phases_ms": phases, "true_relres": true_relres, "true_relres_max": true_relres["poisson_pcg_max"],
                "parity": parity,
                "gpu_launches": launches, "clocks": clocks, "e2e": e2e}
    return cx, line, n


def gpu_same_sample(cx, n, steps, jacobi):
    """The GPU path on the CPU sample's mesh (same n): product defaults, or the CPU port's algorithm (Jacobi-PCG
    everywhere, zero initial guess of the Poisson solve) for a same-algorithm ratio."""
    mv, L, lib, torch = cx.mv, cx.L, cx.lib, cx.torch
    if jacobi:
        L.check(lib.mv_set_mass_solver(L.MV_MASS_PCG))
        L.check(lib.mv_set_poisson_history(0))
    try:
        mesh = mv.UnitCubeMesh(n, n, n, distribute=False)
        mover = mv.MongeAmpereMover(mesh, mv.sinatan3_hessian_monitor(), method="relaxation", rtol=0.0, maxiter=10 ** 6,
                                    dtol=1e300, raise_convergence_errors=False)
        if jacobi:
            mover._dm.set_preconditioner("jacobi")
        mover._begin()
        relax_k(cx, mover, 1, 0)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        _, cg = relax_k(cx, mover, steps, 1)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        return {"value": mesh.num_vertices() * steps / dt, "unit": UNIT, "n": n, "steps": steps,
                "inner_iterations_per_step(l2,poisson,hessian)": [float(x) for x in cg.mean(axis=0)],
                "algorithm": "Jacobi-PCG for every inner solve, zero initial guess of the Poisson solve (the CPU port's algorithm)" if jacobi else
                             "product defaults (Chebyshev mass solves, multigrid PCG, history guess)"}
    finally:
        L.check(lib.mv_set_mass_solver(L.MV_MASS_CHEBYSHEV))
        L.check(lib.mv_set_poisson_history(1))


def run_cfg5(args):
    cx, line, n = run_relaxation(args, 3)
    if cx.rank == 0:
        if cx.world == 1 and not args.no_cpu_baseline:
            cpu_n = args.cpu_n or 96
            r = cpu_relax(cpu_n, args.cpu_steps, 1, args.inner_rtol)
            cb = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample", "n")}
            # the GPU on the very same sample: the like-for-like ratios (same size; same size AND same algorithm)
            cb["gpu_same_sample"] = gpu_same_sample(cx, cpu_n, 5, jacobi=False)
            cb["gpu_same_sample_same_algorithm"] = gpu_same_sample(cx, cpu_n, 3, jacobi=True)
            cb["ratio_same_sample"] = cb["gpu_same_sample"]["value"] / r["value"]
            cb["ratio_same_sample_same_algorithm"] = cb["gpu_same_sample_same_algorithm"]["value"] / r["value"]
            cb["note"] = ("the headline value is measured at n = 256 with multigrid PCG; the CPU port runs Jacobi-PCG on a bounded "
                          f"{cpu_n}^3 sample, so value / cpu_baseline.value mixes size and algorithm -- use the two ratios here")
            line["cpu_baseline"] = cb
        print(json.dumps(line), flush=True)
    cx.parallel.finalize()


def run_cfg4(args):
    if args.method == "quasi_newton":
        return run_cfg4_qn(args)
    cx, line, n = run_relaxation(args, 2)
    if cx.rank == 0:
        if cx.world == 1 and not args.no_cpu_baseline:
            r = cpu_relax(args.cpu_n or 512, args.cpu_steps, 1, args.inner_rtol, dim=2)
            line["cpu_baseline"] = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample", "n")}
        print(json.dumps(line), flush=True)
    cx.parallel.finalize()


def run_cfg4_qn(args):
    """Quasi-Newton on the Helmholtz monitor (demos/monge_ampere_helmholtz.py:134-150): a "step" is one Newton iteration
    (FGMRES solve + l2 line search with its residual evaluations)."""
    cx = Ctx(args)
    np, torch, mv = cx.np, cx.torch, cx.mv
    n = args.n or 8192
    mesh = mv.UnitSquareMesh(n, n)
    Vg = (n + 1) ** 2
    mover = mv.MongeAmpereMover(mesh, mv.helmholtz_hessian_monitor(), method="quasi_newton", rtol=1e-8,
                                maxiter=args.steps, raise_convergence_errors=False)
    sampler = ClockSampler(cx.dev)
    if cx.rank == 0:
        sampler.start()
    cx.barrier()
    t0 = time.perf_counter()
    try:
        its = mover.move()
    except Warning:
        its = mover.snes_iterations
    torch.cuda.synchronize()
    cx.barrier()
    dt = cx.max_over_ranks(time.perf_counter() - t0)
    clocks = sampler.stop() if cx.rank == 0 else None
    if cx.rank == 0:
        its = max(int(its), 1)
        print(json.dumps({"metric": "MA quasi-Newton vertex-iters/s on 2D square mesh (Helmholtz monitor)", "value": Vg * its / dt,
                          "unit": "vertex-Newton-iterations/s", "n_gpus": cx.world, "steps": its, "warmup": 0,
                          "ms_per_step": 1e3 * dt / its, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                          "dtype": "f64", "data": "synthetic",
                          "config": {"workload": workload_name(n, 2) + ", method='quasi_newton'", "n": n, "vertices": Vg,
                                     "ksp_iterations": mover.ksp_iterations, "function_evaluations": mover.function_evaluations,
                                     "gmres_restart_used": getattr(mover, "gmres_restart_used", None),
                                     "residual_trace": [float(t[3]) for t in getattr(mover, "trace", [])]},
                          "clocks": clocks}), flush=True)
    cx.parallel.finalize()


# ---------------------------------------------------------------------------------------------
# configs 2 and 3: forced boundary movement (demos/laplacian_smoothing.py:79-147, demos/lineal_spring.py:99-170)
# ---------------------------------------------------------------------------------------------
def run_forced(args, kind):
    cx = Ctx(args)
    np, torch, mv, L, lib = cx.np, cx.torch, cx.mv, cx.L, cx.lib
    n = args.n or (1024 if kind == "laplacian" else 4096)
    mesh = mv.UnitSquareMesh(n, n)
    Vg = (n + 1) ** 2
    dt_ = 0.1
    nsteps = max(args.steps, 1)
    if kind == "laplacian":
        mover = mv.LaplacianSmoother(mesh, dt_)
    else:
        mover = mv.SpringMover(mesh, dt_, method="lineal")
    top = mv.Function(mover.coord_space)
    moving = mv.DirichletBC(mover.coord_space, top, 4)
    fixed = mv.DirichletBC(mover.coord_space, 0, [1, 2, 3])
    bnodes = torch.as_tensor(moving.nodes, device="cuda")
    period = 1.0

    def update(t):
        # 0.2 sin(2 pi t / T) sin(pi x) on the top boundary (demos/laplacian_smoothing.py:84-90)
        x = mover.mesh.coordinates.tensor[bnodes, 0]
        top.tensor[bnodes, 1] = 0.2 * np.sin(2 * np.pi * t / period) * torch.sin(np.pi * x)
        top.written()

    times = np.arange(0.0, 1.0001, dt_)
    L.check(lib.mv_prof_reset())
    mover.move(times[0], update, [fixed, moving])       # warm-up: hierarchy / workspace set-up
    sampler = ClockSampler(cx.dev)
    if cx.rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    cx.barrier()
    e0.record()
    done = 0
    for k in range(nsteps):
        mover.move(times[1 + k % (len(times) - 1)], update, [fixed, moving])
        done += 1
    e1.record()
    cx.barrier()
    ms = cx.max_over_ranks(e0.elapsed_time(e1))
    clocks = sampler.stop() if cx.rank == 0 else None
    its = mover.cg_iterations[-done:]
    peak, peak_src = peaks()
    dm = mover._dm
    nnzv = dm.nnz / dm.Vown
    if kind == "laplacian":
        per_it = (12 * nnzv + 4 + 104 * 2)       # B_cg(2), SURVEY 8d
    elif Vg >= 4096:
        # assembled spring blocks + block-diagonal multigrid: SpMV over Kxx, Kxy, Kyy (28 B per non-zero + vectors),
        # update / r.z / direction on [2][V] vectors, and one V-cycle per diagonal block (two fp32-valued level-0 passes, P and R,
        # ~30 % more on the coarse levels)
        vcycle = 1.3 * (2 * (8 * nnzv + 36) + 100)
        per_it = (28 * nnzv + 38) + 176 + 2 * vcycle
    else:
        per_it = (4 * nnzv + 16 + 32) + 2 * 88   # matrix-free spring SpMV + the vector kernels of a 2-component CG iteration
    step_bytes = float(np.mean(its)) * per_it * Vg
    gbs = step_bytes / (ms / done * 1e-3) / 1e9
    if cx.rank == 0:
        name = ("LaplacianSmoother + MeshTanglingChecker, forced top-boundary velocity" if kind == "laplacian"
                else "SpringMover lineal, forced top-boundary displacement")
        print(json.dumps({"metric": f"{kind} vertex-timesteps/s on 2D square mesh", "value": Vg * done / (ms * 1e-3),
                          "unit": "vertex-timesteps/s", "n_gpus": cx.world, "steps": done, "warmup": 1, "ms_per_step": ms / done,
                          "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
                          "config": {"workload": f"{name}, UnitSquareMesh({n},{n})" + (" = BASELINE.json configs[%d]" % (1 if kind == "laplacian" else 2)
                                                                                      if n == (1024 if kind == "laplacian" else 4096) else " (reduced size)"),
                                     "n": n, "vertices": Vg, "cg_iterations_per_timestep": [int(i) for i in its],
                                     "vertex_cg_iterations_per_s": Vg * float(np.sum(its)) / (ms * 1e-3),
                                     "volume_ratio": mover.volume_ratio},
                          "roofline": {"bound": "hbm", "kernel": "whole timestep by the SURVEY 8d byte model (CG iterations x bytes per iteration)",
                                       "achieved": gbs, "peak": peak * cx.world, "unit": "GB/s", "frac": gbs / (peak * cx.world),
                                       "peak_source": peak_src, "traffic": None},
                          "gpu_launches": cx.prof()[3], "clocks": clocks}), flush=True)
    cx.parallel.finalize()


def main():
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    elif args.config == 5:
        run_cfg5(args)
    elif args.config == 4:
        run_cfg4(args)
    elif args.config == 2:
        run_forced(args, "laplacian")
    else:
        run_forced(args, "spring")


if __name__ == "__main__":
    main()
