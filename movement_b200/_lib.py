"""ctypes binding of libmovement_b200.so (the C ABI declared in include/movement_b200.h).

There is no CPU fallback: if the shared library is missing, or no CUDA device is visible
when a compute entry point is called, an error is raised.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmovement_b200.so")

MV_OK, MV_ERR_BADARG, MV_ERR_CUDA, MV_ERR_NOT_CONVERGED, MV_ERR_VALENCE, MV_ERR_COMM, MV_ERR_UNSUPPORTED = range(7)
MV_OP_MASS, MV_OP_STIFFNESS, MV_OP_MASS_PLUS_STIFFNESS = 0, 1, 2
MV_PC_JACOBI, MV_PC_AMG = 0, 1
MV_MASS_PCG, MV_MASS_CHEBYSHEV, MV_MASS_CHEBYSHEV_FP64 = 0, 1, 2
MV_MON_CONST, MV_MON_BALL, MV_MON_BALL_RAW, MV_MON_GAUSS_HESS, MV_MON_SINATAN3, MV_MON_USER = 0, 1, 2, 3, 4, 100
(MV_F_XI, MV_F_X, MV_F_PHI, MV_F_PHI_OLD, MV_F_H, MV_F_H_OLD, MV_F_GRAD_PHI, MV_F_MONITOR, MV_F_VOLUME,
 MV_F_ORIG_VOLUME, MV_F_RESIDUAL) = range(11)
(MV_GET_ROW_NNZ, MV_GET_SLICE_PTR, MV_GET_COLS, MV_GET_MASS, MV_GET_STIFF, MV_GET_LUMPED, MV_GET_CELL_VOL,
 MV_GET_BF_NORMAL, MV_GET_BF_AREA, MV_GET_BF_CELL) = range(10)

_vp, _i, _i64, _d = C.c_void_p, C.c_int, C.c_int64, C.c_double
_pi, _pi64, _pd = C.POINTER(C.c_int), C.POINTER(C.c_int64), C.POINTER(C.c_double)

# name -> (restype, argtypes); every symbol declared in include/movement_b200.h
SIGNATURES = {
    "mv_version": (_i, []),
    "mv_last_error": (C.c_char_p, []),
    "mv_device_count": (_i, []),
    "mv_set_device": (_i, [_i]),
    "mv_set_stream": (_i, [_vp]),
    "mv_sync": (_i, []),
    "mv_prof_enable": (_i, [_i]),
    "mv_prof_reset": (_i, []),
    "mv_prof_get": (_i, [_pd, _pi64, _pi64, _pi64]),
    "mv_true_residual_enable": (_i, [_i]),
    "mv_true_residual_get": (_i, [_pd, _pd, _pi64]),
    "mv_mesh_create": (_i, [_i, _i64, _i64, _vp, _vp, _i64, _vp, _vp, _i64, _i64, C.POINTER(_vp)]),
    "mv_mesh_create_ex": (_i, [_i, _i64, _i64, _vp, _vp, _i64, _vp, _vp, _i64, _i64, _i, C.POINTER(_vp)]),
    "mv_mesh_destroy": (None, [_vp]),
    "mv_mesh_info": (_i, [_vp, _pi64]),
    "mv_mesh_scalars": (_i, [_vp, _pd]),
    "mv_mesh_get": (_i, [_vp, _i, _vp]),
    "mv_tangle_init": (_i, [_vp, _vp, _vp]),
    "mv_tangle_check": (_i, [_vp, _vp, _vp, _vp, _pi64]),
    "mv_spmv": (_i, [_vp, _i, _i, _vp, _vp]),
    "mv_pcg_solve": (_i, [_vp, _i, _i, _vp, _vp, _vp, _d, _d, _i, _i, _pi, _pd]),
    "mv_set_preconditioner": (_i, [_vp, _i]),
    "mv_set_mass_solver": (_i, [_i]),
    "mv_amg_info": (_i, [_vp, _i, _pi64, _i, _pd, _pd]),
    "mv_amg_level_get": (_i, [_vp, _i, _i, _i, _pi64, _vp, _vp, _vp, _vp]),
    "mv_ma_create": (_i, [_vp, C.POINTER(_vp)]),
    "mv_ma_destroy": (None, [_vp]),
    "mv_ma_bind": (_i, [_vp, _i, _vp]),
    "mv_recover_gradient_l2": (_i, [_vp, _vp, _vp, _d, _i, _vp]),
    "mv_clement_gradient": (_i, [_vp, _i, _vp, _vp]),
    "mv_project_p1": (_i, [_vp, _vp, _vp, _vp, _d, _i, _pi, _pi64]),
    "mv_ma_set_boundary": (_i, [_vp, _i, _vp]),
    "mv_ma_set_boundary_ex": (_i, [_vp, _i, _vp, _i, _vp, _vp]),
    "mv_ma_get_bcmask": (_i, [_vp, _vp]),
    "mv_ma_set_solver": (_i, [_vp, _d, _d, _i, _i]),
    "mv_ma_set_monitor": (_i, [_vp, _i, _i, _pd]),
    "mv_ma_l2_project": (_i, [_vp, _pi]),
    "mv_ma_update_coordinates": (_i, [_vp]),
    "mv_ma_eval_monitor": (_i, [_vp, _vp, _vp]),
    "mv_ma_diagnostics": (_i, [_vp, _i, _pd]),
    "mv_ma_pseudotimestep": (_i, [_vp, _d, _pi]),
    "mv_set_poisson_history": (_i, [_i]),
    "mv_ma_equidistribute": (_i, [_vp, _pi]),
    "mv_ma_advance": (_i, [_vp]),
    "mv_ma_relax_iterate": (_i, [_vp, _d, _i, _pd, _pi]),
    "mv_ma_relax_move": (_i, [_vp, _d, _i, _i, _i, _d, _d, _pd, _pd, _pi, _pi]),
    "mv_qn_residual": (_i, [_vp, _vp, _pd]),
    "mv_qn_jacobian_apply": (_i, [_vp, _vp, _vp]),
    "mv_qn_solve": (_i, [_vp, _vp, _vp, _d, _i, _i, _d, _i, _pd]),
    "mv_qn_save": (_i, [_vp]),
    "mv_qn_set_iterate": (_i, [_vp, _d, _vp]),
    "mv_ma_l2_rhs": (_i, [_vp, _vp, _vp]),
    "mv_ma_hess_rhs": (_i, [_vp, _vp, _vp]),
    "mv_ma_assemble_a": (_i, [_vp, _vp, _vp, _vp]),
    "mv_cell_volumes": (_i, [_vp, _vp, _vp, _pd]),
    "mv_cell_volumes_static": (_i, [_i, _i64, _vp, _vp, _vp, _pd]),
    "mv_laplace_move": (_i, [_vp, _vp, _vp, _d, _vp, _d, _d, _i, _pi, _pd]),
    "mv_spring_edge_count": (_i, [_vp, _pi64]),
    "mv_spring_edge_geometry": (_i, [_vp, _vp, _vp, _vp, _vp, _vp]),
    "mv_spring_move": (_i, [_vp, _vp, _vp, _vp, _d, _vp, _d, _d, _i, _pi, _pd]),
    "mv_spring_apply": (_i, [_vp, _vp, _vp, _vp, _vp]),
    "mv_spring_solver_info": (_i, [_vp, _pi64]),
    "mv_comm_unique_id": (_i, [C.c_char_p, _vp]),
    "mv_comm_init": (_i, [C.c_char_p, _vp, _i, _i]),
    "mv_comm_finalize": (_i, []),
    "mv_comm_info": (_i, [_pi64]),
    "mv_mesh_set_halo": (_i, [_vp, _i, _vp, _vp, _vp, _vp]),
    "mv_mesh_set_cell_order": (_i, [_vp, _vp]),
    "mv_halo_exchange": (_i, [_vp, _i, _vp]),
    "mv_allreduce_sum": (_i, [_vp, _i]),
}


class MovementB200Error(RuntimeError):
    def __init__(self, status, message):
        super().__init__(f"libmovement_b200 error {status}: {message}")
        self.status = status
        self.message = message


_lib = None


def load():
    """Load the shared library (once).  Raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `python -m movement_b200.build` "
            "(movement_b200 has no CPU fallback)"
        )
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def last_error():
    return load().mv_last_error().decode()


def check(status):
    if status != MV_OK:
        raise MovementB200Error(status, last_error())


def ptr(x):
    """Device/host pointer of a torch tensor or numpy array (None -> NULL)."""
    if x is None:
        return None
    if hasattr(x, "data_ptr"):
        return C.c_void_p(x.data_ptr())
    if hasattr(x, "ctypes"):
        return C.c_void_p(x.ctypes.data)
    if isinstance(x, int):
        return C.c_void_p(x)
    raise TypeError(f"cannot take a pointer of {type(x)}")
