"""ctypes binding of libcfd_b200.so (include/cfd_b200.h).  There is no CPU fallback: a missing library or a
missing GPU raises."""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libcfd_b200.so")

CFD_EQ_COUNT = 8
FAMILY_CAVITY, FAMILY_GENERAL, FAMILY_TEACH = 0, 1, 2
ORDER_REDBLACK, ORDER_SEQUENTIAL = 0, 1

FIELDS = ("u", "v", "p", "pp", "vis", "den", "t", "cp", "ae", "aw", "an", "as", "ap", "su", "sp", "du", "dv",
          "gamt", "susrc", "spsrc", "te", "ed", "gen", "gamh")
FIELD_ID = {n: i for i, n in enumerate(FIELDS)}
GEOMS = ("x", "xu", "sew", "sewu", "y", "yv", "sns", "snsv",
         "xdif", "xcvi", "xcvip", "fx", "fxm",
         "ydif", "ycvr", "ycvrs", "arx", "arxj", "arxjp", "r", "rmn", "sx", "sxmn", "fv", "fvp", "fy", "fym",
         "dxep", "dxpw", "dxepu", "dxpwu", "dynp", "dyps", "dynpv", "dypsv", "rv", "rcv")
GEOM_ID = {n: i for i, n in enumerate(GEOMS)}


class CfdConfig(C.Structure):
    _fields_ = [("family", C.c_int), ("ni", C.c_int), ("nj", C.c_int), ("iadv", C.c_int), ("order", C.c_int),
                ("nswp", C.c_int * CFD_EQ_COUNT), ("urf", C.c_double * CFD_EQ_COUNT),
                ("ipref", C.c_int), ("jpref", C.c_int), ("device", C.c_int), ("tdma_m", C.c_int),
                ("goff", C.c_int), ("ni_global", C.c_int), ("halo_w", C.c_int), ("halo_e", C.c_int),
                ("sew_nim1", C.c_double), ("solve_flow", C.c_int), ("solve_t", C.c_int),
                ("indcos", C.c_int), ("istep", C.c_int), ("jstep", C.c_int), ("_pad", C.c_int),
                ("viscos", C.c_double), ("flowin", C.c_double), ("great", C.c_double),
                ("joff", C.c_int), ("nj_global", C.c_int), ("halo_s", C.c_int), ("halo_n", C.c_int),
                ("sns_njm1", C.c_double), ("raw_coefficients", C.c_int)]


class CfdHookDesc(C.Structure):
    _fields_ = [("kind", C.c_int), ("params", C.c_double * 4)]


HOOK_NONE, HOOK_BOUSSINESQ, HOOK_FIELD_SOURCES = 0, 1, 2
EQ_U, EQ_V, EQ_P, EQ_T, EQ_TE, EQ_ED = 0, 1, 2, 3, 4, 5
WALL_ID = {"taun": 0, "tauw": 1, "yplusn": 2, "xplusw": 3}


class CfdTurbulence(C.Structure):
    _fields_ = ([(n, C.c_double) for n in ("cmu", "cd", "c1", "c2", "cappa", "elog", "pred", "prte", "prandl", "prandt",
                                           "pfun", "twalln", "urfk", "urfe", "urft", "urfvis")]
                + [(n, C.c_int) for n in ("nswpk", "nswpd", "nswpt", "incalk", "incald", "incalt", "inpro")])


class CfdHeatConfig(C.Structure):
    _fields_ = [("nx", C.c_int), ("ny", C.c_int), ("k", C.c_double), ("rho", C.c_double), ("cp", C.c_double),
                ("xa", C.c_double), ("xb", C.c_double), ("ya", C.c_double), ("yb", C.c_double),
                ("tleft", C.c_double), ("tright", C.c_double), ("tbottom", C.c_double), ("ttop", C.c_double),
                ("dt", C.c_double), ("qthr", C.c_double), ("qval", C.c_double),
                ("device", C.c_int), ("j_goff", C.c_int), ("ny_global", C.c_int),
                ("halo_s", C.c_int), ("halo_n", C.c_int)]


class CfdFdcavConfig(C.Structure):
    _fields_ = [("ni", C.c_int), ("nj", C.c_int), ("re", C.c_double), ("dt", C.c_double), ("utop", C.c_double),
                ("urfp", C.c_double), ("device", C.c_int), ("steady", C.c_int), ("urfu", C.c_double), ("urfv", C.c_double)]


FDC_FIELDS = ("u", "v", "p", "uo", "vo", "po", "pp", "su", "ae", "aw", "an", "as", "ap")


class CfdError(RuntimeError):
    pass


# every symbol include/cfd_b200.h declares: name -> (restype, argtypes)
_vp, _dp, _ip = C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_int)
SYMBOLS = {
    "cfd_create": (C.c_int, [C.POINTER(_vp), C.POINTER(CfdConfig)]),
    "cfd_destroy": (C.c_int, [_vp]),
    "cfd_strerror": (C.c_char_p, [C.c_int]),
    "cfd_last_error": (C.c_char_p, []),
    "cfd_set_stream": (C.c_int, [_vp, _vp]),
    "cfd_synchronize": (C.c_int, [_vp]),
    "cfd_set_geometry": (C.c_int, [_vp, C.c_int, _vp, C.c_int]),
    "cfd_upload": (C.c_int, [_vp, C.c_int, _vp, C.c_int]),
    "cfd_download": (C.c_int, [_vp, C.c_int, _vp, C.c_int]),
    "cfd_upload_noalloc": (C.c_int, [_vp, C.c_int, _vp, C.c_int]),
    "cfd_probe": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int, _dp]),
    "cfd_extract_line": (C.c_int, [_vp, C.c_int, C.c_int, _dp, C.c_int]),
    "cfd_extract_row": (C.c_int, [_vp, C.c_int, C.c_int, _dp, C.c_int]),
    "cfd_centreline": (C.c_int, [_vp, C.c_int, C.c_int, _dp, C.c_int]),
    "cfd_reference_pressure": (C.c_int, [_vp]),
    "cfd_field_device_view": (C.c_int, [_vp, C.c_int, C.POINTER(_vp), C.POINTER(C.c_longlong), _ip]),
    "cfd_layout_info": (C.c_int, [_vp, _ip, _ip, _ip]),
    "cfd_launch_count": (C.c_longlong, [_vp]),
    "cfd_calcu": (C.c_int, [_vp, _dp]),
    "cfd_calcv": (C.c_int, [_vp, _dp]),
    "cfd_calcp": (C.c_int, [_vp, _dp]),
    "cfd_calct": (C.c_int, [_vp, _dp]),
    "cfd_set_hook": (C.c_int, [_vp, C.c_int, C.POINTER(CfdHookDesc)]),
    "cfd_assemble_t": (C.c_int, [_vp]),
    "cfd_lisolv": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int]),
    "cfd_iterate": (C.c_int, [_vp, C.c_int, _dp]),
    "cfd_assemble_u": (C.c_int, [_vp]),
    "cfd_assemble_v": (C.c_int, [_vp]),
    "cfd_assemble_p": (C.c_int, [_vp]),
    "cfd_sweep": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]),
    "cfd_sweep_part": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]),
    "cfd_correct": (C.c_int, [_vp, _vp]),
    "cfd_sweeps": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int]),
    "cfd_ipc_export": (C.c_int, [_vp, C.c_int, _vp]),
    "cfd_ipc_open": (C.c_int, [_vp, _vp, C.POINTER(_vp)]),
    "cfd_peer_attach": (C.c_int, [_vp, C.c_int, C.c_int, _vp, C.c_int, C.c_int, C.c_int]),
    "cfd_peer_detach": (C.c_int, [_vp]),
    "cfd_halo_push": (C.c_int, [_vp, C.c_int, _ip, _ip, _ip]),
    "cfd_peer_status": (C.c_int, [_vp, _ip]),
    "cfd_bcast_attach": (C.c_int, [_vp, C.c_int, _vp, C.c_int]),
    "cfd_scalar_bcast": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int, C.c_int, C.POINTER(_vp)]),
    "cfd_sweeps_single_launch": (C.c_int, [_vp, _ip]),
    "cfd_debug_flow_timing": (C.c_int, [_vp, _dp]),
    "cfd_set_turbulence": (C.c_int, [_vp, C.POINTER(CfdTurbulence)]),
    "cfd_wall_upload": (C.c_int, [_vp, C.c_int, _vp, C.c_int]),
    "cfd_wall_download": (C.c_int, [_vp, C.c_int, _vp, C.c_int]),
    "cfd_calcte": (C.c_int, [_vp, _dp]),
    "cfd_calced": (C.c_int, [_vp, _dp]),
    "cfd_props": (C.c_int, [_vp]),
    "cfd_assemble_te": (C.c_int, [_vp]),
    "cfd_assemble_ed": (C.c_int, [_vp]),
    "cfd_lin_wesweep": (C.c_int, [_vp, C.c_int]),
    "cfd_lin_snsweep": (C.c_int, [_vp, C.c_int]),
    "cfd_lin_residual": (C.c_int, [_vp, C.c_int, _dp]),
    "cfd_lin_solve": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double, _ip, _dp]),
    "cfd_lin_time_terms": (C.c_int, [_vp, C.c_int, C.c_int, C.POINTER(C.c_int), C.c_int]),
    "cfd_field_copy": (C.c_int, [_vp, C.c_int, C.c_int]),
    "cfd_spike_basis": (C.c_int, [_vp, C.c_int, C.c_int]),
    "cfd_spike_pack": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, _vp]),
    "cfd_spike_finish": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int, C.c_int, _vp, _vp, C.c_int, C.c_int]),
    "cfd_rows_pack": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int, _vp]),
    "cfd_rows_unpack": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int, _vp]),
    "cfd_residuals": (C.c_int, [_vp, _dp]),
    "cfd_residual_device_ptr": (C.c_int, [_vp, C.POINTER(_vp)]),
    "cfd_heat_create": (C.c_int, [C.POINTER(_vp), C.POINTER(CfdHeatConfig)]),
    "cfd_heat_destroy": (C.c_int, [_vp]),
    "cfd_heat_set_stream": (C.c_int, [_vp, _vp]),
    "cfd_heat_upload": (C.c_int, [_vp, _vp, C.c_int]),
    "cfd_heat_download": (C.c_int, [_vp, _vp, C.c_int]),
    "cfd_heat2d_step": (C.c_int, [_vp, C.c_int]),
    "cfd_heat2d_step_exact": (C.c_int, [_vp, C.c_int]),
    "cfd_heat2d_implicit_step": (C.c_int, [_vp, C.c_int]),
    "cfd_heat2d_implicit_step_exact": (C.c_int, [_vp, C.c_int]),
    "cfd_heat_implicit_halo": (C.c_int, [C.c_double, C.POINTER(C.c_int)]),
    "cfd_sipsol": (C.c_int, [C.c_int, C.c_int, C.c_int, C.c_int] + [_vp] * 7 + [C.c_int, C.c_double, C.c_double, C.c_int,
                             C.POINTER(C.c_int), C.POINTER(C.c_double)]),
    "cfd_heat_set_operator": (C.c_int, [_vp, _vp, _vp, _vp, _vp, _vp, C.c_int]),
    "cfd_heat2d_apply": (C.c_int, [_vp, _vp, C.c_int]),
    "cfd_heat2d_step_var": (C.c_int, [_vp, C.c_int]),
    "cfd_heat2d_steady_gs": (C.c_int, [_vp, C.c_int, C.c_double, C.c_double, _ip, _dp]),
    "cfd_heat_device_view": (C.c_int, [_vp, C.POINTER(_vp), C.POINTER(C.c_longlong), _ip]),
    "cfd_heat_synchronize": (C.c_int, [_vp]),
    "cfd_fdcav_create": (C.c_int, [C.POINTER(_vp), C.POINTER(CfdFdcavConfig)]),
    "cfd_fdcav_destroy": (C.c_int, [_vp]),
    "cfd_fdcav_upload": (C.c_int, [_vp, C.c_int, _vp, C.c_int]),
    "cfd_fdcav_download": (C.c_int, [_vp, C.c_int, _vp, C.c_int]),
    "cfd_fdcav_predict": (C.c_int, [_vp]),
    "cfd_fdcav_correct": (C.c_int, [_vp, _dp]),
    "cfd_fdcav_solve_pp": (C.c_int, [_vp, _ip]),
    "cfd_fdcav_step": (C.c_int, [_vp, _dp, _ip]),
    "cfd_fdcav_implicit_assemble": (C.c_int, [_vp, C.c_int]),
    "cfd_fdcav_implicit_solve": (C.c_int, [_vp, C.c_int, _ip]),
    "cfd_fdcav_implicit_correct": (C.c_int, [_vp, _dp]),
    "cfd_fdcav_implicit_step": (C.c_int, [_vp, _dp, _ip]),
    "cfd_version": (C.c_char_p, []),
}

_lib = None


def load() -> C.CDLL:
    """dlopen the CUDA library; raises if it has not been built (python __graft_entry__.py build)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise CfdError(f"{LIB_PATH} is missing: build it with `make -C cfd_b200/csrc` "
                           "(there is no CPU fallback)")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def check(rc: int) -> None:
    if rc != 0:
        lib = load()
        raise CfdError(f"{lib.cfd_strerror(rc).decode()}: {lib.cfd_last_error().decode()}")
