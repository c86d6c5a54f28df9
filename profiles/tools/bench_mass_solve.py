"""Micro-benchmark of the mass solves at full size: the mixed-precision Chebyshev solver (default), the all-double
one and the Jacobi-PCG, cold start, rtol 1e-12.  Launch-shape variants through MV_CHEB_VAR (one process each).

    python profiles/tools/bench_mass_solve.py [n] [ncomp,ncomp,...] [kinds: mixed,fp64,pcg]
"""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from movement_b200 import _lib as L
from movement_b200.core import DeviceMesh
from movement_b200.mesh import cube_arrays
n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
comps = [int(c) for c in sys.argv[2].split(",")] if len(sys.argv) > 2 else [9, 3, 1]
kinds = sys.argv[3].split(",") if len(sys.argv) > 3 else ["mixed", "fp64", "pcg"]
KIND = {"mixed": L.MV_MASS_CHEBYSHEV, "fp64": L.MV_MASS_CHEBYSHEV_FP64, "pcg": L.MV_MASS_PCG}
lib = L.load()
dm = DeviceMesh(*cube_arrays(n, n, n))
V = dm.Vown
g = torch.Generator(device="cuda").manual_seed(1)
var = os.environ.get("MV_CHEB_VAR", "0")
for ncomp in comps:
    b = torch.randn(V, ncomp, dtype=torch.float64, device="cuda", generator=g) / V
    if ncomp == 1:
        b = b[:, 0].contiguous()
    res = {}
    for name in kinds:
        L.check(lib.mv_set_mass_solver(KIND[name]))
        best = 1e30
        for rep in range(4):
            x = torch.zeros_like(b)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            it, rel = dm.pcg(L.MV_OP_MASS, b, x, rtol=1e-12)
            e1.record()
            torch.cuda.synchronize()
            if rep:
                best = min(best, e0.elapsed_time(e1))
        res[name] = x
        print(f"var={var} n={n} c={ncomp} {name}: it={it} rel={rel.max():.2e} {best:.2f} ms/solve  {best / max(it, 1):.3f} ms/it", flush=True)
    ref = res.get("pcg", res.get("fp64"))
    for name in kinds:
        if ref is not None and res[name] is not ref:
            print(f"   {name} vs {'pcg' if 'pcg' in res else 'fp64'}: max rel diff {float((res[name] - ref).abs().max() / ref.abs().max()):.2e}")
L.check(lib.mv_set_mass_solver(L.MV_MASS_CHEBYSHEV))
