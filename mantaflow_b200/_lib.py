"""ctypes binding of libmantacuda.so (include/mantacuda.h).

The product path has no CPU fallback: importing this module without the built CUDA library raises,
and creating a solver without a CUDA device raises (mf_solver_create fails in the library).
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libmantacuda.so")


class MantaError(RuntimeError):
    """Mirrors Manta::Error -> Python RuntimeError (source/general.h:42-57, pwrapper/pclass.cpp:57-61)."""


class PressureParams(C.Structure):
    _fields_ = [
        ("cgAccuracy", C.c_float), ("phi", C.c_void_p), ("perCellCorr", C.c_void_p), ("fractions", C.c_void_p),
        ("obvel", C.c_void_p), ("gfClamp", C.c_float), ("cgMaxIterFac", C.c_float), ("precondition", C.c_int),
        ("preconditioner", C.c_int), ("enforceCompatibility", C.c_int), ("useL2Norm", C.c_int),
        ("zeroPressureFixing", C.c_int), ("curv", C.c_void_p), ("surfTens", C.c_float), ("retRhs", C.c_void_p),
    ]


class ShapeDesc(C.Structure):          # mf_shape
    _fields_ = [("type", C.c_int), ("a", C.c_float * 3), ("b", C.c_float * 3), ("r", C.c_float)]


class NoiseParams(C.Structure):        # mf_noise_params
    _fields_ = [("posOffset", C.c_float * 3), ("posScale", C.c_float * 3), ("valOffset", C.c_float), ("valScale", C.c_float),
                ("clamp", C.c_int), ("clampNeg", C.c_float), ("clampPos", C.c_float), ("timeAnim", C.c_float)]


class SolveInfo(C.Structure):
    _fields_ = [("iterations", C.c_int), ("resNorm", C.c_float), ("maxIter", C.c_int), ("fixedCell", C.c_int),
                ("mgLevels", C.c_int), ("mgRebuilt", C.c_int)]


# every symbol declared in include/mantacuda.h: name -> (restype, argtypes)
_P = C.c_void_p
_f = C.c_float
_i = C.c_int
_s = C.c_char_p
SIGNATURES = {
    "mf_version": (_i, []),
    "mf_last_error": (_s, []),
    "mf_device_count": (_i, []),
    "mf_build_info": (_s, []),
    "mf_solver_create": (_P, [_i, _i, _i, _i, _i]),
    "mf_solver_destroy": (None, [_P]),
    "mf_solver_set_dt": (_i, [_P, _f]),
    "mf_solver_get_dt": (_f, [_P]),
    "mf_solver_sync": (_i, [_P]),
    "mf_solver_stream": (_P, [_P]),
    "mf_solver_launch_count": (C.c_longlong, [_P]),
    "mf_solver_pool_bytes": (C.c_longlong, [_P]),
    "mf_solver_profile_begin": (_i, [_P]),
    "mf_solver_profile_end": (_i, [_P, C.c_char_p, _i]),
    "mf_solver_event_record": (_i, [_P, _i]),
    "mf_solver_event_elapsed_ms": (_i, [_P, _i, _i, C.POINTER(_f)]),
    "mf_comm_create": (_P, [_i, _i, _i, _s]),
    "mf_comm_create_emulated": (_P, [_i, _i]),
    "mf_comm_destroy": (None, [_P]),
    "mf_solver_create_dist": (_P, [_P, _i, _i, _i, _i]),
    "mf_solver_ranks": (_i, [_P]),
    "mf_solver_slab": (_i, [_P, _i, C.POINTER(_i), C.POINTER(_i)]),
    "mf_solver_barrier_count": (C.c_longlong, [_P]),
    "mf_grid_upload_slab": (_i, [_P, _P]),
    "mf_grid_download_slab": (_i, [_P, _P]),
    "mf_grid_stage_upload": (_i, [_P, _P]),
    "mf_grid_commit_upload": (_i, [_P]),
    "mf_grid_download_async": (_i, [_P, _P]),
    "mf_solver_transfers_sync": (_i, [_P]),
    "mf_comm_selftest": (_i, [_P, C.c_longlong, _i, C.POINTER(C.c_double)]),
    "mf_host_alloc": (_P, [C.c_longlong]),
    "mf_host_free": (None, [_P]),
    "mf_grid_create": (_P, [_P, _i]),
    "mf_grid_destroy": (None, [_P]),
    "mf_grid_type": (_i, [_P]),
    "mf_grid_clear": (_i, [_P]),
    "mf_grid_copy": (_i, [_P, _P]),
    "mf_grid_swap": (_i, [_P, _P]),
    "mf_grid_get_component": (_i, [_P, _P, _i]),
    "mf_grid_set_component": (_i, [_P, _P, _i]),
    "mf_grid_upload": (_i, [_P, _P]),
    "mf_grid_download": (_i, [_P, _P]),
    "mf_grid_set_const_real": (_i, [_P, _f]),
    "mf_grid_set_const_int": (_i, [_P, _i]),
    "mf_grid_set_const_vec3": (_i, [_P, _f, _f, _f]),
    "mf_grid_add_scaled": (_i, [_P, _P, _f, _f, _f]),
    "mf_grid_mult_const": (_i, [_P, _f, _f, _f]),
    "mf_grid_max_abs": (_i, [_P, C.POINTER(_f)]),
    "mf_grid_min_max": (_i, [_P, C.POINTER(_f), C.POINTER(_f)]),
    "mf_grid_sum_sqr": (_i, [_P, C.POINTER(C.c_double)]),
    "mf_grid_dot": (_i, [_P, _P, C.POINTER(C.c_double)]),
    "mf_grid_max_diff": (_i, [_P, _P, C.POINTER(C.c_double)]),
    "mf_dot_product_noalloc": (_i, [_P, _P, C.c_longlong, _P, _P]),
    "mf_flags_init_domain": (_i, [_P, _i, _s, _s, _s, _s, _P]),
    "mf_flags_fill_grid": (_i, [_P, _i]),
    "mf_flags_count_cells": (_i, [_P, _i, _i, C.POINTER(_i)]),
    "mf_flags_set_open_bound": (_i, [_P, _i, _s, _i]),
    "mf_apply_box_to_real": (_i, [_P, _f, _f, _f, _f, _f, _f, _f]),
    "mf_apply_box_to_flags": (_i, [_P, _f, _f, _f, _f, _f, _f, _i]),
    "mf_apply_cylinder_to_real": (_i, [_P, _f, _f, _f, _f, _f, _f, _f, _f]),
    "mf_advect_semi_lagrange": (_i, [_P, _P, _P, _i, _f, _i, _i, _i]),
    "mf_set_wall_bcs": (_i, [_P, _P, _P, _P, _P, _i]),
    "mf_add_buoyancy": (_i, [_P, _P, _P, _f, _f, _f, _f, _i]),
    "mf_add_gravity": (_i, [_P, _P, _f, _f, _f, _P, _i]),
    "mf_set_inflow_bcs": (_i, [_P, _s, _f, _f, _f]),
    "mf_reset_outflow": (_i, [_P, _P, _P]),
    "mf_pressure_params_default": (None, [C.POINTER(PressureParams)]),
    "mf_compute_pressure_rhs": (_i, [_P, _P, _P, _P, C.POINTER(PressureParams)]),
    "mf_solve_pressure_system": (_i, [_P, _P, _P, _P, C.POINTER(PressureParams), C.POINTER(SolveInfo)]),
    "mf_correct_velocity": (_i, [_P, _P, _P, C.POINTER(PressureParams)]),
    "mf_solve_pressure": (_i, [_P, _P, _P, C.POINTER(PressureParams), C.POINTER(SolveInfo)]),
    "mf_release_mg": (_i, [_P]),
    "mf_shape_apply": (_i, [_P, _P, _f, _f, _f, _i, _P]),
    "mf_shape_levelset": (_i, [_P, _P]),
    "mf_flags_update_from_levelset": (_i, [_P, _P]),
    "mf_levelset_join": (_i, [_P, _P]),
    "mf_update_fractions": (_i, [_P, _P, _P, _i, _f]),
    "mf_set_obstacle_flags": (_i, [_P, _P, _P, _P, _P, _i]),
    "mf_parts_create": (_P, [_P]),
    "mf_parts_destroy": (None, [_P]),
    "mf_parts_size": (C.c_longlong, [_P]),
    "mf_parts_create_pdata": (_i, [_P, _i, C.POINTER(_i)]),
    "mf_parts_upload": (_i, [_P, C.POINTER(_f), C.POINTER(_i), C.c_longlong]),
    "mf_parts_download": (_i, [_P, C.POINTER(_f), C.POINTER(_i)]),
    "mf_pdata_upload": (_i, [_P, _i, _P]),
    "mf_pdata_download": (_i, [_P, _i, _P]),
    "mf_sample_flags_with_particles": (_i, [_P, _P, _i, _f]),
    "mf_parts_advect_in_grid": (_i, [_P, _P, _P, _i, _i, _i, _i]),
    "mf_map_parts_to_mac": (_i, [_P, _P, _P, _P, _i, _P]),
    "mf_mark_fluid_cells": (_i, [_P, _P, _P]),
    "mf_map_mac_to_parts": (_i, [_P, _P, _P, _i]),
    "mf_flip_velocity_update": (_i, [_P, _P, _P, _P, _i, _f]),
    "mf_noise_create": (_P, [_P, _i, _i]),
    "mf_noise_destroy": (None, [_P]),
    "mf_noise_generate_tile": (_i, [C.POINTER(_f)]),
    "mf_noise_seed_offset": (_i, [_P, C.POINTER(_f)]),
    "mf_noise_evaluate": (_i, [_P, C.POINTER(NoiseParams), _f, C.POINTER(_f), _i, C.POINTER(_f)]),
    "mf_density_inflow": (_i, [_P, _P, _P, C.POINTER(NoiseParams), _P, _f, _f, _f]),
    "mf_add_noise": (_i, [_P, _P, _P, C.POINTER(NoiseParams), _P, _f, _f]),
    "mf_extrapolate_mac_simple": (_i, [_P, _P, _i, _P, _i]),
    "mf_extrapolate_mac_from_weight": (_i, [_P, _P, _i]),
    "mf_extrapolate_ls_simple": (_i, [_P, _i, _i, _i]),
    "mf_extrapolate_vec3_simple": (_i, [_P, _P, _i, _i]),
    "mf_make_laplace_matrix": (_i, [_P, _P, _P, _P, _P, _P]),
    "mf_apply_matrix": (_i, [_P, _P, _P, _P, _P, _P, _P]),
    "mf_apply_laplace_rowcode": (_i, [_P, _P, _P]),
    "mf_cg_solve_diffusion": (_i, [_P, _P, _f, _f, _f]),
    "mf_cg_create": (_P, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "mf_cg_destroy": (None, [_P]),
    "mf_cg_set_ic_preconditioner": (_i, [_P, _i]),
    "mf_cg_set_mg_preconditioner": (_i, [_P, _i, _i]),
    "mf_cg_set_accuracy": (_i, [_P, _f]),
    "mf_cg_set_use_l2_norm": (_i, [_P, _i]),
    "mf_cg_force_reinit": (_i, [_P]),
    "mf_cg_iterate": (_i, [_P, C.POINTER(_i)]),
    "mf_cg_solve": (_i, [_P, _i]),
    "mf_cg_get": (_i, [_P, C.POINTER(_f), C.POINTER(_f), C.POINTER(_f), C.POINTER(_f)]),
    "mf_mg_info": (_i, [_P, C.POINTER(_i), C.POINTER(_i)]),
    "mf_mg_level_size": (_i, [_P, _i, C.POINTER(_i)]),
    "mf_mg_level_format": (_i, [_P, _i, C.POINTER(_i)]),
    "mf_mg_download_level": (_i, [_P, _i, _P, _P]),
    "mf_mg_vcycle": (_i, [_P, _P, _P, _P, _P, _P, _P, _f]),
}

_lib = None


def lib():
    """Load libmantacuda.so once; fail loudly if it was not built (no fallback path exists)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                "mantaflow_b200: %s is missing -- build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(there is no CPU fallback)" % LIB_PATH)
        l = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(l, name)   # AttributeError if the library does not export a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = l
    return _lib


def check(status):
    if status != 0:
        msg = lib().mf_last_error()
        raise MantaError(msg.decode("utf-8", "replace") if msg else "mantacuda error %d" % status)
