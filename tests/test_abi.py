"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol that
include/movement_b200.h declares, the ctypes table covers them, and the product path refuses to
run without a GPU (no CPU fallback, nothing under movement_b200/ touches oracle/)."""
import ctypes as C
import os
import re

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "movement_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mv_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    from movement_b200 import _lib as L

    lib = L.load()
    names = _declared()
    assert len(names) >= 45
    for n in names:
        assert hasattr(lib, n), f"{n} declared in movement_b200.h but not exported"
        assert n in L.SIGNATURES, f"{n} has no ctypes signature"
    assert set(L.SIGNATURES) <= set(names), "ctypes table binds undeclared symbols"
    assert lib.mv_version() >= 100


def test_no_cpu_fallback():
    from movement_b200 import _lib as L

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    lib = L.load()
    assert lib.mv_device_count() == 0
    coords = np.array([[0.0, 0.0], [1.0, 0.0], [0.0, 1.0]])
    cells = np.array([[0, 1, 2]], dtype=np.int32)
    h = C.c_void_p()
    st = lib.mv_mesh_create(2, 3, 1, L.ptr(coords), L.ptr(cells), 0, None, None, 3, 1, C.byref(h))
    assert st == L.MV_ERR_CUDA and b"no CPU fallback" in lib.mv_last_error()
    import movement_b200 as mv

    with pytest.raises(RuntimeError, match="no CPU fallback"):
        mv.MongeAmpereMover(mv.UnitSquareMesh(2, 2), mv.ConstantMonitorBuilder(2)())
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        mv.LaplacianSmoother(mv.UnitSquareMesh(2, 2), 1.0)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "movement_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f), encoding="utf-8").read()
                assert "movement_oracle" not in text and "oracle_c" not in text, f


def test_host_mesh_surface_without_gpu():
    import movement_b200 as mv

    mesh = mv.UnitCubeMesh(3, 3, 3)
    assert mesh.num_vertices() == 64 and mesh.num_cells() == 162 and mesh.topological_dimension == 3
    assert mesh.exterior_facets.unique_markers == [1, 2, 3, 4, 5, 6]
    bc = mv.DirichletBC(mesh.coordinates.function_space(), 0, "on_boundary")
    assert len(bc.nodes) == 64 - 8
    f = mv.Function(mesh.coordinates.function_space()).interpolate([1.0, 0.0, 0.0])
    assert np.allclose(f.dat.data[:, 0], 1.0)
    copy = mv.Mesh(mesh.coordinates.copy(deepcopy=True))
    copy.coordinates.dat.data[:] += 1.0
    assert np.allclose(copy.coordinates.dat.data - mesh.coordinates.dat.data, 1.0)
    u = mv.Function(mesh.function_space("P1"))
    for b in (mv.GradientMonitorBuilder, mv.HessianMonitorBuilder):
        with pytest.raises(AssertionError):          # movement/monitor.py:188: the solution must be a Function
            b(3, None, 1.0)
        with pytest.raises(AssertionError):          # :225 / :278: positive scale factor
            b(3, u, 0.0)
        assert callable(b(3, u, 1.0).get_monitor())
    spec = mv.RingMonitorBuilder((0.5, 0.5), 0.4, 20.0, 200.0)()(mesh)
    assert spec.params == (0.5, 0.5, 0.0, 0.4, 20.0, 200.0)


def test_vtk_output_and_profiling_events(tmp_path):
    """f4: the demos' VTKFile(...).write(mesh.coordinates) and the reference's PETSc log events as named ranges."""
    import xml.etree.ElementTree as ET

    import movement_b200 as mv
    import movement_b200.profiling as prof

    mesh = mv.UnitSquareMesh(3, 2)
    out = mv.VTKFile(str(tmp_path / "run.pvd"))
    phi = mv.Function(mesh.function_space("P1"), name="phi").interpolate(lambda x: x[:, 0] + 2 * x[:, 1])
    for t in (0.0, 0.5):
        out.write(mesh.coordinates, phi, time=t)
    sets = ET.parse(tmp_path / "run.pvd").getroot().find("Collection").findall("DataSet")
    assert [s.get("timestep") for s in sets] == ["0", "0.5"] and sets[1].get("file") == "run_1.vtu"
    piece = ET.parse(tmp_path / "run_1.vtu").getroot().find("UnstructuredGrid").find("Piece")
    assert piece.get("NumberOfPoints") == "12" and piece.get("NumberOfCells") == "12"
    names = [a.get("Name") for a in piece.find("PointData").findall("DataArray")]
    assert names == ["Coordinates", "phi"]
    vals = np.array(piece.find("PointData").findall("DataArray")[1].text.split(), dtype=float)
    assert np.allclose(vals, phi.dat.data_ro)
    prof.reset()
    prof.enable()
    try:
        with prof.event("movement.monge_ampere.MongeAmpereMover_Relaxation.move"):
            pass
        assert prof.events()["movement.monge_ampere.MongeAmpereMover_Relaxation.move"][0] == 1
        assert "MongeAmpereMover_Relaxation.move" in prof.log_view()
    finally:
        prof.enable(False)
    # every site the reference decorates is decorated here
    for cls, name in ((mv.MongeAmpereMover_Relaxation, "move"), (mv.MongeAmpereMover_QuasiNewton, "move"), (mv.LaplacianSmoother, "move"),
                      (mv.SpringMover_Lineal, "move"), (mv.MeshTanglingChecker, "check")):
        assert hasattr(getattr(cls, name), "__wrapped__")


def test_vtk_output_of_a_partitioned_mesh_is_one_piece_per_rank(tmp_path, monkeypatch):
    """every rank used to write the same out_<k>.vtu with its ghosts.  Now: out_<k>_<rank>.vtu with the owned
    cells only, the .pvd (rank 0) lists all pieces; the pieces together hold every cell of the global mesh once."""
    import xml.etree.ElementTree as ET

    import movement_b200 as mv
    from movement_b200 import parallel
    from movement_b200.mesh import Mesh, square_arrays
    from movement_b200.partition import partition_mesh

    arrays = square_arrays(6, 5)
    total, ghosts = 0, 0
    for rank in (1, 0):                       # rank 0 last: it writes the index
        monkeypatch.setattr(parallel, "_state", {"active": True, "rank": rank, "size": 2})
        lp = partition_mesh(*arrays, rank, 2)
        mesh = Mesh(lp.coords, lp.cells, lp.bfacets, lp.btags, part=lp, device="cpu")
        vol = mv.Function(mesh.function_space("P0"), name="volume")
        vol.dat.data[:] = np.arange(mesh.num_cells())
        out = mv.VTKFile(str(tmp_path / "run.pvd"))
        path = out.write(mesh.coordinates, vol, time=0.25)
        assert path.endswith(f"run_0_{rank}.vtu")
        piece = ET.parse(path).getroot().find("UnstructuredGrid").find("Piece")
        assert int(piece.get("NumberOfCells")) == mesh.n_owned_cells <= mesh.num_cells()
        ghosts += mesh.num_cells() - mesh.n_owned_cells
        cd = np.array(piece.find("CellData").find("DataArray").text.split(), dtype=float)
        assert np.array_equal(cd, np.arange(mesh.n_owned_cells))
        total += mesh.n_owned_cells
    assert total == arrays[1].shape[0] and ghosts > 0
    sets = ET.parse(tmp_path / "run.pvd").getroot().find("Collection").findall("DataSet")
    assert [(s.get("part"), s.get("file")) for s in sets] == [("0", "run_0_0.vtu"), ("1", "run_0_1.vtu")]
