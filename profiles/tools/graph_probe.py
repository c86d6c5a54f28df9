import sys, os
sys.path.insert(0, "/root/repo")
import torch, movement_b200 as mv
mv.mover._print = mv.monge_ampere._print = lambda *a, **k: None
mesh = mv.UnitCubeMesh(48, 48, 48)
m = mv.MongeAmpereMover(mesh, mv.sinatan3_hessian_monitor(), method="relaxation", rtol=1e-3, maxiter=8, raise_convergence_errors=False)
try:
    m.move()
except Warning:
    pass
print("cg", m.cg_iterations[:4])
