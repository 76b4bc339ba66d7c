"""pyi_fill_loess_static: the allocation-free variant of fill::loess (scratch in
a static device array, include/pyinterp_b200.h) against pyi_fill_loess, both
called through the C-ABI on device pointers.  The bar is the one of the loess
tests (tests/test_gpu_fill.py): identical bits.

Collected last on purpose; it passes alone in
a fresh process and after every other test."""
import ctypes as C
import importlib

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _pod(abi, first_guess, periodic, iterations, hx, hy, value_type, epsilon):
    return abi.LoessConfig(first_guess, int(periodic), iterations, hx, hy, value_type, epsilon)


def _field(shape, dtype, seed):
    rng = np.random.default_rng(seed)
    x = np.linspace(0.0, 4.0 * np.pi, shape[0])[:, None]
    y = np.linspace(-1.0, 1.0, shape[1])[None, :]
    g = (np.sin(x) * (1.0 + y * y) + 0.1 * rng.standard_normal(shape)).astype(dtype)
    g[rng.random(shape) < 0.2] = np.nan
    g[shape[0] // 4: shape[0] // 4 + 9, shape[1] // 3: shape[1] // 3 + 11] = np.nan
    return g


def test_static_scratch_loess_matches_the_allocating_entry_point(pyi):
    import torch
    abi = importlib.import_module("pangeo-pyinterp_b200._abi")
    lib = abi.load()
    limit = 8 << 20  # PYI_LOESS_STATIC_MAX_BYTES
    stream = torch.cuda.current_stream().cuda_stream
    # two calls in a row on the shared array: the second on a SMALLER grid of the
    # other dtype, so anything left behind by the first call would be seen
    calls = [
        # (shape, dtype, first guess, periodic, iterations, hx, hy, value type, epsilon)
        ((1024, 1000), "float64", 1, True, 6, 3, 3, 0, 1e-9),   # 8 192 000 bytes: close to the limit
        ((1440, 720), "float32", 0, False, 4, 2, 5, 0, 0.0),    # the 0.25 degree grid
        ((1440, 720), "float32", 1, True, 30, 4, 2, 2, 1e-3),   # stops on epsilon, every cell processed
        ((37, 23), "float64", 1, True, 1, 3, 3, 1, 1e-4),       # single pass (no scratch at all)
        ((64, 5), "float64", 1, False, 0, 3, 3, 0, 1e-4),       # first guess only
    ]
    for k, (shape, dtype, fg, per, its, hx, hy, vt, eps) in enumerate(calls):
        g = _field(shape, dtype, 900 + k)
        assert g.nbytes <= limit
        src = torch.from_numpy(g).cuda()
        want = torch.full_like(src, -7.0)
        got = torch.full_like(src, -9.0)
        pod = _pod(abi, fg, per, its, hx, hy, vt, eps)
        dt = abi.DTYPES[dtype]
        abi.check(lib.pyi_fill_loess(src.data_ptr(), want.data_ptr(), dt, shape[0], shape[1], C.byref(pod),
                                     abi.PYI_MEM_DEVICE, stream), "loess")
        abi.check(lib.pyi_fill_loess_static(src.data_ptr(), got.data_ptr(), dt, shape[0], shape[1],
                                            C.byref(pod), stream), "loess (static scratch)")
        torch.cuda.synchronize()
        a, b = got.cpu().numpy(), want.cpu().numpy()
        assert np.array_equal(src.cpu().numpy(), g, equal_nan=True)  # the input is left untouched
        assert np.array_equal(a, b, equal_nan=True), (shape, dtype, fg, per, its)
        if vt == 0 and its != 1:
            assert not np.isnan(a).any()  # iterated from a first guess: every hole is filled
        # the same call again: same bits (nothing carried over in the array)
        got2 = torch.full_like(src, -11.0)
        abi.check(lib.pyi_fill_loess_static(src.data_ptr(), got2.data_ptr(), dt, shape[0], shape[1],
                                            C.byref(pod), stream), "loess (static scratch)")
        assert np.array_equal(got2.cpu().numpy(), b, equal_nan=True)


def test_static_scratch_loess_rejects_grids_beyond_its_limit(pyi):
    import torch
    abi = importlib.import_module("pangeo-pyinterp_b200._abi")
    lib = abi.load()
    pod = _pod(abi, 1, True, 3, 3, 3, 0, 1e-4)
    src = torch.zeros((1025, 1024), dtype=torch.float64, device="cuda")  # 8 396 800 bytes > 8 MiB
    out = torch.empty_like(src)
    st = lib.pyi_fill_loess_static(src.data_ptr(), out.data_ptr(), abi.DTYPES["float64"], 1025, 1024,
                                   C.byref(pod), None)
    assert st == abi.PYI_BAD_SHAPE
    assert b"exceeds the static scratch of 8388608 bytes" in lib.pyi_last_error()
    with pytest.raises(ValueError, match="exceeds the static scratch"):
        abi.check(st, "loess (static scratch)")
    # exactly at the limit: accepted
    src = torch.zeros((1024, 1024), dtype=torch.float64, device="cuda")
    src[::7, ::5] = float("nan")
    out = torch.empty_like(src)
    st = lib.pyi_fill_loess_static(src.data_ptr(), out.data_ptr(), abi.DTYPES["float64"], 1024, 1024,
                                   C.byref(pod), None)
    assert st == abi.PYI_OK
    assert torch.all(out == 0).item()
    # the usual argument errors keep the reference's text
    bad = _pod(abi, 1, True, 3, 0, 3, 0, 1e-4)
    st = lib.pyi_fill_loess_static(src.data_ptr(), out.data_ptr(), abi.DTYPES["float64"], 1024, 1024,
                                   C.byref(bad), None)
    assert st == abi.PYI_BAD_ARG and b"at least 1" in lib.pyi_last_error()
