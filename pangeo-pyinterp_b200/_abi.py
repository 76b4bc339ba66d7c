"""ctypes mirror of ``include/pyinterp_b200.h`` and loader of the CUDA library.

The shared library ``csrc/libpyinterp_b200.so`` is the product: hand-written
sm_100a kernels behind a C-ABI.  There is no CPU fallback -- if the library is
missing, loading fails loudly; if there is no CUDA device, every compute entry
point returns ``PYI_NO_DEVICE`` and the Python layer raises.
"""
from __future__ import annotations

import ctypes as C
import os

# ---- status codes / enums (include/pyinterp_b200.h) -------------------------
PYI_OK, PYI_OOB, PYI_BAD_SHAPE, PYI_BAD_ARG, PYI_NO_DEVICE, PYI_CUDA_ERROR, PYI_NOMEM, PYI_OVERFLOW = range(8)
PYI_AXIS_REGULAR, PYI_AXIS_IRREGULAR = 0, 1
PYI_COORD_F64, PYI_COORD_I64 = 0, 1
PYI_BILINEAR, PYI_IDW, PYI_NEAREST = 0, 1, 2
PYI_AXIS_LINEAR, PYI_AXIS_NEAREST = 0, 1
PYI_W_BICUBIC, PYI_W_BILINEAR = 0, 1
#: pyi_spatial_method: every fitting model of the windowed path
SPATIAL_METHODS = {
    "bicubic": 0, "bilinear": 1, "akima": 2, "akima_periodic": 3, "c_spline": 4,
    "c_spline_not_a_knot": 5, "c_spline_periodic": 6, "linear": 7, "steffen": 8, "polynomial": 9,
}
#: numpy datetime unit -> pyi_resolution (dateutils::DType::Resolution)
RESOLUTIONS = {"Y": 0, "M": 1, "W": 2, "D": 3, "h": 4, "m": 5, "s": 6, "ms": 7, "us": 8, "ns": 9,
               "ps": 10, "fs": 11, "as": 12}
#: pyi_first_guess (config::fill::FirstGuess, config/fill.hpp:18-23)
FIRST_GUESS = {"zonal_average": 0, "zero": 1}
PYI_BOUNDARY_UNDEF, PYI_BOUNDARY_SHRINK = 0, 1
PYI_MEM_HOST, PYI_MEM_DEVICE = 0, 1
PYI_VALUES_DEFAULT, PYI_VALUES_EXACT, PYI_VALUES_TOLERANCE = 0, 1, 2
VALUE_MODES = {"default": 0, "exact": 1, "tolerance": 2}
PYI_GRID_DEFAULT, PYI_GRID_PACK_CELLS, PYI_GRID_BORROW_DEVICE = 0, 1, 2

#: numpy dtype name -> pyi_dtype (pybind/grid.hpp:604-606,819-879 of the reference)
DTYPES = {
    "int8": 0, "uint8": 1, "int16": 2, "uint16": 3, "int32": 4,
    "uint32": 5, "int64": 6, "uint64": 7, "float32": 8, "float64": 9,
}


def result_dtype(grid_dtype: str) -> str:
    """Result type rule of pybind/grid_dispatch.hpp:25-30."""
    return "float64" if grid_dtype in ("float64", "int64", "uint64") else "float32"


class Scalar(C.Union):
    _fields_ = [("f", C.c_double), ("i", C.c_int64)]


class AxisDesc(C.Structure):
    _fields_ = [
        ("kind", C.c_int32),
        ("coord_type", C.c_int32),
        ("size", C.c_int64),
        ("is_ascending", C.c_int32),
        ("is_periodic", C.c_int32),
        ("period", Scalar),
        ("start", Scalar),
        ("step", Scalar),
        ("min_value", Scalar),
        ("max_value", Scalar),
        ("inv_step", C.c_double),
        ("step_2", C.c_int64),
        ("points", C.c_void_p),
    ]


class GeometricConfig(C.Structure):
    _fields_ = [
        ("method", C.c_int32),
        ("idw_exp", C.c_int32),
        ("z_method", C.c_int32),
        ("u_method", C.c_int32),
        ("bounds_error", C.c_int32),
        ("num_threads", C.c_int32),
        ("value_mode", C.c_int32),
    ]


class WindowedConfig(C.Structure):
    _fields_ = [
        ("method", C.c_int32),
        ("half_x", C.c_int32),
        ("half_y", C.c_int32),
        ("boundary", C.c_int32),
        ("bounds_error", C.c_int32),
        ("num_threads", C.c_int32),
        ("value_mode", C.c_int32),
    ]


class WindowedNdConfig(C.Structure):
    _fields_ = [
        ("method", C.c_int32),
        ("half_x", C.c_int32),
        ("half_y", C.c_int32),
        ("boundary", C.c_int32),
        ("bounds_error", C.c_int32),
        ("num_threads", C.c_int32),
        ("z_method", C.c_int32),
        ("u_method", C.c_int32),
        ("derivative", C.c_int32),
        ("value_mode", C.c_int32),
    ]


class FillConfig(C.Structure):
    _fields_ = [
        ("first_guess", C.c_int32),
        ("is_periodic", C.c_int32),
        ("max_iterations", C.c_uint64),
        ("epsilon", C.c_double),
        ("relaxation", C.c_double),
        ("pre_smooth", C.c_uint32),
        ("post_smooth", C.c_uint32),
    ]


class LoessConfig(C.Structure):
    _fields_ = [
        ("first_guess", C.c_int32),
        ("is_periodic", C.c_int32),
        ("max_iterations", C.c_uint32),
        ("half_x", C.c_uint32),
        ("half_y", C.c_uint32),
        ("value_type", C.c_int32),
        ("epsilon", C.c_double),
    ]


#: every symbol include/pyinterp_b200.h declares: name -> (restype, argtypes)
_P = C.c_void_p
_I64P = C.POINTER(C.c_int64)
_IP = C.POINTER(C.c_int)
SIGNATURES = {
    "pyi_abi_version": (C.c_int, []),
    "pyi_device_count": (C.c_int, []),
    "pyi_set_device": (C.c_int, [C.c_int]),
    "pyi_last_error": (C.c_char_p, []),
    "pyi_dist_init": (C.c_int, [C.c_int]),
    "pyi_dist_size": (C.c_int, []),
    "pyi_axis_create_f64": (_P, [_P, C.c_int64, C.c_double, C.c_int, C.c_double, _IP]),
    "pyi_axis_create_i64": (_P, [_P, C.c_int64, C.c_int64, C.c_int, C.c_int64, _IP]),
    "pyi_axis_create_from_desc": (_P, [C.POINTER(AxisDesc), _IP]),
    "pyi_axis_flip": (C.c_int, [_P]),
    "pyi_axis_describe": (C.c_int, [_P, C.POINTER(AxisDesc)]),
    "pyi_axis_destroy": (None, [_P]),
    "pyi_axis_find_indexes": (C.c_int, [_P, C.c_int64, _P, _P, C.c_int, _P]),
    "pyi_axis_find_index": (C.c_int, [_P, C.c_int64, _P, C.c_int, _P, C.c_int, _P]),
    "pyi_grid_create": (_P, [C.c_int, C.POINTER(_P), C.c_int, _P, _I64P, C.c_int, C.c_uint, _IP]),
    "pyi_grid_update": (C.c_int, [_P, _P, C.c_int]),
    "pyi_grid_pack": (C.c_int, [_P]),
    "pyi_grid_track_host": (C.c_int, [_P, _P]),
    "pyi_grid_mark_stale": (C.c_int, [_P]),
    "pyi_grid_destroy": (None, [_P]),
    "pyi_geometric": (C.c_int, [_P, C.POINTER(GeometricConfig), C.c_int64, _P, _P, _P, _P, _P,
                                C.c_int, _I64P, _IP, _P]),
    "pyi_windowed2d": (C.c_int, [_P, C.POINTER(WindowedConfig), C.c_int64, _P, _P, _P, C.c_int,
                                 _I64P, _IP, _P]),
    "pyi_windowed": (C.c_int, [_P, C.POINTER(WindowedNdConfig), C.c_int64, _P, _P, _P, _P, _P, C.c_int,
                               _I64P, _IP, _P]),
    "pyi_datetime_convert": (C.c_int, [C.c_int, C.c_int, C.c_int64, _P, _P, C.c_int, _I64P, _IP, _P]),
    "pyi_fill_gauss_seidel": (C.c_int, [_P, C.c_int, C.c_int64, C.c_int64, C.POINTER(FillConfig), C.c_int,
                                        C.POINTER(C.c_uint64), C.POINTER(C.c_double), _P]),
    "pyi_fill_multigrid": (C.c_int, [_P, C.c_int, C.c_int64, C.c_int64, C.POINTER(FillConfig), C.c_int,
                                     C.POINTER(C.c_uint64), C.POINTER(C.c_double), _P]),
    "pyi_fill_loess": (C.c_int, [_P, _P, C.c_int, C.c_int64, C.c_int64, C.POINTER(LoessConfig), C.c_int, _P]),
    "pyi_fill_loess_static": (C.c_int, [_P, _P, C.c_int, C.c_int64, C.c_int64, C.POINTER(LoessConfig), _P]),
    "pyi_format_bounds_error": (C.c_int, [_P, C.c_int, _P, C.c_char_p, C.c_size_t]),
    "pyi_format_temporal": (C.c_int, [C.c_int, C.c_int, C.c_int64, C.c_char_p, C.c_size_t]),
    "pyi_model_stencil": (C.c_int, [C.c_int, C.c_double, _P]),
    "pyi_launch_count": (C.c_int64, []),
}

ABI_VERSION = 2
# PYINTERP_B200_LIB: another build of the same library (csrc/Makefile `check`: bounds-checked gathers)
LIB_PATH = os.environ.get("PYINTERP_B200_LIB") or os.path.join(
    os.path.dirname(os.path.abspath(__file__)), "csrc", "libpyinterp_b200.so")
_lib = None


def bind(lib, signatures=SIGNATURES):
    """Attach restype/argtypes; raises AttributeError on a missing symbol."""
    for name, (res, args) in signatures.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    return lib


def load():
    """Load the CUDA library.  No fallback: a missing build is an error."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python __graft_entry__.py build` "
                "(nvcc, sm_100a).  pyinterp_b200 has no CPU fallback."
            )
        _lib = bind(C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL))
        if _lib.pyi_abi_version() != ABI_VERSION:
            raise ImportError("libpyinterp_b200.so: ABI version mismatch")
    return _lib


class PyiError(RuntimeError):
    """A C-ABI call failed (CUDA error, no device, bad argument)."""


def check(status: int, what: str = "") -> None:
    if status in (PYI_OK, PYI_OOB):
        return
    msg = load().pyi_last_error()
    text = msg.decode() if msg else ""
    names = {PYI_BAD_SHAPE: "BAD_SHAPE", PYI_BAD_ARG: "BAD_ARG", PYI_NO_DEVICE: "NO_DEVICE",
             PYI_CUDA_ERROR: "CUDA_ERROR", PYI_NOMEM: "NOMEM"}
    if status == PYI_OVERFLOW:
        raise OverflowError(text)  # std::overflow_error (dateutils.hpp:55,99)
    if status in (PYI_BAD_SHAPE, PYI_BAD_ARG):
        raise ValueError(f"{what}: {text or names[status]}")
    raise PyiError(f"{what}: {names.get(status, status)}: {text}")
