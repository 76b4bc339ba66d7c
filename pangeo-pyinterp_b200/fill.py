"""Filling of undefined grid values, the mirror of ``pyinterp.fill``
(pyinterp/fill.py) for the methods that run on the device:

* :func:`gauss_seidel` -- ``core.fill.gauss_seidel``
  (cxx/include/pyinterp/fill/gauss_seidel.hpp:166-211);
* :func:`multigrid` -- ``core.fill.multigrid``
  (cxx/include/pyinterp/fill/multi_grid.hpp:276-334);
* :func:`loess` -- ``core.fill.loess``
  (cxx/include/pyinterp/fill/loess.hpp:207-253), which returns a new array.

Grids are filled IN PLACE like the reference: numpy arrays are staged through
the device, CUDA tensors (torch, or anything exposing
``__cuda_array_interface__`` / a CUDA ``__dlpack__``) are relaxed where they
live.  Results are bit-identical to the reference build (black cells first,
see csrc/pyi_fill.cu).  There is no CPU fallback.
"""
from __future__ import annotations

import ctypes as C
from typing import Any

import numpy as np

from . import _abi
from .config import fill as _cfg
from .core import _device_view, _is_torch_cuda, _lib, _np_ptr

__all__ = ["gauss_seidel", "multigrid", "loess"]

#: pyinterp/fill.py:47-50
_FIRST_GUESS_MAP = {
    "zero": _cfg.FirstGuess.ZERO,
    "zonal_average": _cfg.FirstGuess.ZONAL_AVERAGE,
}


#: pyinterp/fill.py:53-57
_LOESS_VALUE_TYPE_MAP = {
    "all": _cfg.LoessValueType.ALL,
    "defined": _cfg.LoessValueType.DEFINED,
    "undefined": _cfg.LoessValueType.UNDEFINED,
}


def _build_config(config_class, method_map, **kwargs):
    """pyinterp/fill.py:78-106: keyword arguments -> configuration object."""
    cfg = config_class()
    for name, value in kwargs.items():
        if value is None or name not in method_map:
            continue
        method_name, value_map = method_map[name]
        cfg = getattr(cfg, method_name)(value_map[value] if value_map else value)
    return cfg


def _grid_args(grid: Any):
    """(array or tensor, pointer, dtype code, nx, ny, mem, stream)."""
    grid = _device_view(grid)
    if _is_torch_cuda(grid):
        import torch
        if grid.dim() != 2 or grid.dtype not in (torch.float32, torch.float64):
            raise TypeError("grid must be a 2-D float32 or float64 array")
        if not grid.is_contiguous():
            raise ValueError("device grids must be C-contiguous (filled in place)")
        dt = _abi.DTYPES["float64" if grid.dtype == torch.float64 else "float32"]
        return (grid, grid.data_ptr(), dt, grid.shape[0], grid.shape[1], _abi.PYI_MEM_DEVICE,
                torch.cuda.current_stream(grid.device).cuda_stream)
    if not isinstance(grid, np.ndarray) or grid.ndim != 2 or grid.dtype not in (np.float32, np.float64):
        # nanobind reports an overload mismatch as TypeError
        raise TypeError("grid must be a 2-D float32 or float64 array")
    if not grid.flags.writeable:
        raise TypeError("grid must be writeable (filled in place)")
    return grid, None, _abi.DTYPES[grid.dtype.name], grid.shape[0], grid.shape[1], _abi.PYI_MEM_HOST, None


def gauss_seidel(grid, config=None, *, max_iterations: int | None = None, epsilon: float | None = None,
                 relaxation: float | None = None, first_guess: str | None = None,
                 is_periodic: bool | None = None, num_threads: int | None = None) -> tuple[int, float]:
    """``pyinterp.fill.gauss_seidel`` (pyinterp/fill.py:181-237): replaces the
    NaN cells of ``grid`` in place by the red-black Gauss-Seidel solution of the
    Laplace equation; returns ``(iterations, max_residual)``."""
    if config is None:
        config = _build_config(
            _cfg.GaussSeidel,
            {"max_iterations": ("with_max_iterations", None), "epsilon": ("with_epsilon", None),
             "relaxation": ("with_relaxation", None), "first_guess": ("with_first_guess", _FIRST_GUESS_MAP),
             "is_periodic": ("with_is_periodic", None), "num_threads": ("with_num_threads", None)},
            max_iterations=max_iterations, epsilon=epsilon, relaxation=relaxation, first_guess=first_guess,
            is_periodic=is_periodic, num_threads=num_threads)
    elif not isinstance(config, _cfg.GaussSeidel):
        raise TypeError("config must be a config.fill.GaussSeidel")
    pod = _abi.FillConfig(int(config.first_guess()), int(config.is_periodic()), config.max_iterations(),
                          config.epsilon(), config.relaxation(), 0, 0)
    return _run("pyi_fill_gauss_seidel", "gauss_seidel", grid, pod)


def _run(entry: str, what: str, grid, pod) -> tuple[int, float]:
    arr, ptr, dt, nx, ny, mem, stream = _grid_args(grid)
    its, res = C.c_uint64(0), C.c_double(0.0)
    fn = getattr(_lib(), entry)
    if mem == _abi.PYI_MEM_DEVICE:
        import torch
        with torch.cuda.device(arr.device):
            st = fn(ptr, dt, nx, ny, C.byref(pod), mem, C.byref(its), C.byref(res), stream)
    else:
        # the reference accepts any strided view (EigenDRef): fill a dense copy
        # and write it back when the caller's array is not C-contiguous
        dense = arr if arr.flags.c_contiguous else np.ascontiguousarray(arr)
        st = fn(_np_ptr(dense), dt, nx, ny, C.byref(pod), mem, C.byref(its), C.byref(res), None)
        if dense is not arr and st == _abi.PYI_OK:
            arr[...] = dense
    _abi.check(st, what)
    return int(its.value), float(res.value)


def multigrid(grid, config=None, *, max_iterations: int | None = None, epsilon: float | None = None,
              pre_smooth: int | None = None, post_smooth: int | None = None, first_guess: str | None = None,
              is_periodic: bool | None = None, num_threads: int | None = None) -> tuple[int, float]:
    """``pyinterp.fill.multigrid`` (pyinterp/fill.py:308-372): replaces the NaN
    cells of ``grid`` in place by V-cycle multigrid relaxation of the Laplace
    equation; returns ``(iterations, max_residual)``."""
    if config is None:
        config = _build_config(
            _cfg.Multigrid,
            {"max_iterations": ("with_max_iterations", None), "epsilon": ("with_epsilon", None),
             "pre_smooth": ("with_pre_smooth", None), "post_smooth": ("with_post_smooth", None),
             "first_guess": ("with_first_guess", _FIRST_GUESS_MAP), "is_periodic": ("with_is_periodic", None),
             "num_threads": ("with_num_threads", None)},
            max_iterations=max_iterations, epsilon=epsilon, pre_smooth=pre_smooth, post_smooth=post_smooth,
            first_guess=first_guess, is_periodic=is_periodic, num_threads=num_threads)
    elif not isinstance(config, _cfg.Multigrid):
        raise TypeError("config must be a config.fill.Multigrid")
    pod = _abi.FillConfig(int(config.first_guess()), int(config.is_periodic()), config.max_iterations(),
                          config.epsilon(), 1.0, config.pre_smooth(), config.post_smooth())
    return _run("pyi_fill_multigrid", "multigrid", grid, pod)


def loess(data, config=None, *, nx: int | None = None, ny: int | None = None, max_iterations: int | None = None,
          epsilon: float | None = None, value_type: str | None = None, first_guess: str | None = None,
          is_periodic: bool | None = None, num_threads: int | None = None):
    """``pyinterp.fill.loess`` (pyinterp/fill.py:240-305): a NEW array in which
    every processed cell (``value_type``: the undefined ones, the defined ones,
    or all) is the tri-cube weighted mean of the defined values in its
    ``(2 nx + 1) x (2 ny + 1)`` window; ``max_iterations == 1`` is the single
    pass, larger values iterate from the first guess until the largest change is
    below ``epsilon``.  numpy in, numpy out; a CUDA tensor gives a CUDA tensor."""
    if config is None:
        config = _build_config(
            _cfg.Loess,
            {"nx": ("with_nx", None), "ny": ("with_ny", None), "max_iterations": ("with_max_iterations", None),
             "epsilon": ("with_epsilon", None), "value_type": ("with_value_type", _LOESS_VALUE_TYPE_MAP),
             "first_guess": ("with_first_guess", _FIRST_GUESS_MAP), "is_periodic": ("with_is_periodic", None),
             "num_threads": ("with_num_threads", None)},
            nx=nx, ny=ny, max_iterations=max_iterations, epsilon=epsilon, value_type=value_type,
            first_guess=first_guess, is_periodic=is_periodic, num_threads=num_threads)
    elif not isinstance(config, _cfg.Loess):
        raise TypeError("config must be a config.fill.Loess")
    if config.max_iterations() >= 2 ** 32:
        raise TypeError("max_iterations must fit an uint32")
    pod = _abi.LoessConfig(int(config.first_guess()), int(config.is_periodic()), int(config.max_iterations()),
                           config.nx(), config.ny(), int(config.value_type()), config.epsilon())
    data = _device_view(data)
    lib = _lib()
    if _is_torch_cuda(data):
        import torch
        if data.dim() != 2 or data.dtype not in (torch.float32, torch.float64):
            raise TypeError("data must be a 2-D float32 or float64 array")
        src = data.contiguous()
        out = torch.empty_like(src)
        dt = _abi.DTYPES["float64" if src.dtype == torch.float64 else "float32"]
        with torch.cuda.device(src.device):
            st = lib.pyi_fill_loess(src.data_ptr(), out.data_ptr(), dt, src.shape[0], src.shape[1], C.byref(pod),
                                    _abi.PYI_MEM_DEVICE, torch.cuda.current_stream(src.device).cuda_stream)
        _abi.check(st, "loess")
        return out
    if not isinstance(data, np.ndarray) or data.ndim != 2 or data.dtype not in (np.float32, np.float64):
        raise TypeError("data must be a 2-D float32 or float64 array")
    src = np.ascontiguousarray(data)
    out = np.empty_like(src)
    st = lib.pyi_fill_loess(_np_ptr(src), _np_ptr(out), _abi.DTYPES[src.dtype.name], src.shape[0], src.shape[1],
                            C.byref(pod), _abi.PYI_MEM_HOST, None)
    _abi.check(st, "loess")
    return out
