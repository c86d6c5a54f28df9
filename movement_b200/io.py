"""Mesh / field output for ParaView, standing in for ``firedrake.VTKFile`` as the reference's demos use it
(demos/laplacian_smoothing.py, demos/lineal_spring.py: ``VTKFile("...pvd").write(mesh.coordinates)``).

``VTKFile("out.pvd").write(f, g, time=t)`` writes one ``out_<k>.vtu`` (XML UnstructuredGrid, ASCII) per call with the
current coordinates of the functions' mesh and the functions as point (P1) or cell (P0) data, and keeps ``out.pvd``
(the time-series index) up to date.  Host-side and small: fields are read through ``Function.dat.data_ro``.

On a partitioned mesh (``movement_b200.parallel`` active) every rank writes its own piece ``out_<k>_<rank>.vtu`` holding
the cells it OWNS (ghost cells would be drawn twice) with the local vertices they use, and rank 0 alone writes the
``.pvd``, one ``<DataSet part="rank">`` per piece -- what Firedrake's parallel VTK output amounts to for a reader.
"""
from __future__ import annotations

import os

import numpy as np

_VTK_CELL = {1: 3, 2: 5, 3: 10}   # line, triangle, tetrahedron


class VTKFile:
    def __init__(self, filename, mode="w"):
        if not filename.endswith(".pvd"):
            raise ValueError("VTKFile: the file name must end in .pvd")
        self.filename = filename
        self._base = filename[:-4]
        self._entries = []
        if mode == "a" and os.path.exists(filename):
            raise NotImplementedError("VTKFile: append mode")

    @staticmethod
    def _array(f, name, data, ncomp):
        f.write(f'<DataArray type="Float64" Name="{name}" NumberOfComponents="{ncomp}" format="ascii">\n')
        np.savetxt(f, data.reshape(-1, ncomp), fmt="%.17g")
        f.write("</DataArray>\n")

    def write(self, *functions, time=None):
        from .mesh import Function, Mesh

        if len(functions) == 1 and isinstance(functions[0], Mesh):
            mesh, functions = functions[0], ()
        else:
            assert functions and all(isinstance(f, Function) for f in functions)
            mesh = functions[0].space.mesh
        if getattr(mesh, "periodic", False):
            raise NotImplementedError("VTKFile: periodic meshes (DG coordinates)")
        d = mesh.geometric_dimension
        xyz = np.zeros((mesh.num_vertices(), 3))
        xyz[:, :d] = np.asarray(mesh.coordinates.dat.data_ro).reshape(mesh.num_vertices(), d)
        cells = np.asarray(mesh.cells)
        k = len(self._entries)
        part = getattr(mesh, "part", None)
        rank, nparts = 0, 1
        if part is not None:
            from . import parallel

            rank, nparts = parallel.rank(), parallel.size()
            cells = cells[:mesh.n_owned_cells]        # owned cells come first in the local numbering
        ncell_out = cells.shape[0]
        path = f"{self._base}_{k}_{rank}.vtu" if part is not None else f"{self._base}_{k}.vtu"
        with open(path, "w", encoding="utf-8") as f:
            f.write('<?xml version="1.0"?>\n<VTKFile type="UnstructuredGrid" version="0.1" byte_order="LittleEndian">\n')
            f.write(f'<UnstructuredGrid>\n<Piece NumberOfPoints="{xyz.shape[0]}" NumberOfCells="{cells.shape[0]}">\n<Points>\n')
            self._array(f, "Points", xyz, 3)
            f.write("</Points>\n<Cells>\n")
            f.write('<DataArray type="Int64" Name="connectivity" format="ascii">\n')
            np.savetxt(f, cells, fmt="%d")
            f.write('</DataArray>\n<DataArray type="Int64" Name="offsets" format="ascii">\n')
            np.savetxt(f, (np.arange(cells.shape[0]) + 1) * cells.shape[1], fmt="%d")
            f.write('</DataArray>\n<DataArray type="UInt8" Name="types" format="ascii">\n')
            np.savetxt(f, np.full(cells.shape[0], _VTK_CELL[mesh.topological_dimension]), fmt="%d")
            f.write("</DataArray>\n</Cells>\n")
            point = [g for g in functions if g.space.kind in ("P1", "P1_vec", "P1_ten")]
            cell = [g for g in functions if g.space.kind == "P0"]
            if len(point) + len(cell) != len(functions):
                raise NotImplementedError("VTKFile: only P1 (scalar, vector, tensor) and P0 fields")
            f.write("<PointData>\n")
            for g in point:
                a = np.asarray(g.dat.data_ro)
                if g.space.kind == "P1_vec":          # ParaView wants 3-vectors
                    a3 = np.zeros((a.shape[0], 3)); a3[:, :d] = a; a = a3
                self._array(f, g.name or "function", a,
                            1 if g.space.kind == "P1" else (3 if g.space.kind == "P1_vec" else d * d))
            f.write("</PointData>\n<CellData>\n")
            for g in cell:
                self._array(f, g.name or "function",
                            np.asarray(g.dat.data_ro)[:ncell_out], 1)
            f.write("</CellData>\n</Piece>\n</UnstructuredGrid>\n</VTKFile>\n")
        base = os.path.basename(self._base)
        names = [f"{base}_{k}_{r}.vtu" for r in range(nparts)] if part is not None else [os.path.basename(path)]
        self._entries.append((float(k) if time is None else float(time), names))
        if rank == 0:                                  # the other ranks' piece names are known without asking them
            with open(self.filename, "w", encoding="utf-8") as f:
                f.write('<?xml version="1.0"?>\n<VTKFile type="Collection" version="0.1" byte_order="LittleEndian">\n<Collection>\n')
                for t, pieces in self._entries:
                    for r, name in enumerate(pieces):
                        f.write(f'<DataSet timestep="{t:.17g}" part="{r}" file="{name}"/>\n')
                f.write("</Collection>\n</VTKFile>\n")
        return path
