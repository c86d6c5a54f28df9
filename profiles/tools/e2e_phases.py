import os, sys, time
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import movement_b200 as mv
from movement_b200.mesh import Mesh, cube_arrays
mv.mover._print = lambda *a, **k: None
mv.monge_ampere._print = mv.mover._print
n = int(sys.argv[1]) if len(sys.argv) > 1 else 256
arrays = cube_arrays(n, n, n)
pinned = [torch.from_numpy(np.ascontiguousarray(a)).pin_memory() for a in arrays]
host = tuple(t.numpy() for t in pinned)
def T():
    torch.cuda.synchronize(); return time.perf_counter()
for rep in range(2):
    t0 = T()
    mesh = Mesh(*host)
    t1 = T()
    mover = mv.MongeAmpereMover(mesh, mv.sinatan3_hessian_monitor(), method="relaxation", rtol=1e-8, maxiter=10**6, dtol=1e300)
    t2 = T()
    mover._begin()
    t3 = T()
    mover._evaluate()
    t4 = T()
    mover._advance_potential()
    t5 = T()
    mover._evaluate(); mover._advance_potential()
    t6 = T()
    mover.to_physical_coordinates()
    out = mover.mesh.coordinates.tensor.cpu()
    t7 = T()
    buf = torch.empty(mover.mesh.coordinates.tensor.shape, dtype=torch.float64, pin_memory=True)
    t8 = T()
    buf.copy_(mover.mesh.coordinates.tensor, non_blocking=True)
    t9 = T()
    print(f"rep {rep}: Mesh() {t1-t0:.3f}  Mover() {t2-t1:.3f}  _begin {t3-t2:.3f}  first evaluate {t4-t3:.3f}  first advance (AMG set-up + solves) {t5-t4:.3f}  "
          f"second step {t6-t5:.3f}  to_physical + .cpu() {t7-t6:.3f}  pinned alloc {t8-t7:.3f}  pinned D2H {t9-t8:.3f}", flush=True)
    info = mover._dm.amg_info()
    print("   amg set-up ms", info and info["setup_ms"], "levels", info and info["rows"])
    del mover, mesh, out, buf
    torch.cuda.empty_cache()
