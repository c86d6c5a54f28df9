"""Micro-benchmark of the mass solves at full size (scratch; variants through MV_CHEB_CG)."""
import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from movement_b200 import _lib as L
from movement_b200.core import DeviceMesh
from movement_b200.mesh import cube_arrays
n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
lib = L.load()
dm = DeviceMesh(*cube_arrays(n, n, n))
V = dm.Vown
g = torch.Generator(device="cuda").manual_seed(1)
for ncomp in (9, 3, 1):
    b = torch.randn(V, ncomp, dtype=torch.float64, device="cuda", generator=g) / V
    if ncomp == 1: b = b[:, 0].contiguous()
    res = {}
    for kind, name in ((L.MV_MASS_CHEBYSHEV, "cheb"), (L.MV_MASS_PCG, "pcg")):
        L.check(lib.mv_set_mass_solver(kind))
        for rep in range(3):
            x = torch.zeros_like(b)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            it, rel = dm.pcg(L.MV_OP_MASS, b, x, rtol=1e-12)
            e1.record(); torch.cuda.synchronize()
            ms = e0.elapsed_time(e1)
        res[name] = x
        nnz = dm.nnz
        byt = (12 * nnz + (12 + 32 * ncomp) * V) if name == "cheb" else (12 * nnz + (4 + 104 * ncomp) * V)
        print(f"n={n} c={ncomp} {name}: it={it} rel={rel.max():.2e} {ms:.2f} ms  {ms/max(it,1):.3f} ms/it  {byt*it/ms/1e6:.0f} GB/s(model)", flush=True)
    print("   max diff", float((res["cheb"] - res["pcg"]).abs().max() / res["pcg"].abs().max()))
