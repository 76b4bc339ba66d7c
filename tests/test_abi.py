"""C-ABI boundary checks that need no GPU: the product library loads, exports
every entry point ``include/pyinterp_b200.h`` declares, its host-side handles
(axes) work, and every compute entry point refuses to run without a CUDA
device instead of falling back to the CPU."""
import ctypes as C
import importlib
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "pyinterp_b200.h")
LIB = os.path.join(ROOT, "pangeo-pyinterp_b200", "csrc", "libpyinterp_b200.so")


def _declared():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    names = re.findall(r"\b(pyi_[a-z0-9_]+)\s*\(", text)
    return sorted(set(names))


def test_header_declares_the_path():
    names = _declared()
    for must in ("pyi_geometric", "pyi_windowed2d", "pyi_axis_find_indexes", "pyi_grid_create",
                 "pyi_axis_create_from_desc", "pyi_format_bounds_error"):
        assert must in names


def test_library_exports_every_declared_symbol():
    assert os.path.exists(LIB), "csrc/libpyinterp_b200.so is not built (python __graft_entry__.py)"
    lib = C.CDLL(LIB)
    missing = [n for n in _declared() if not hasattr(lib, n)]
    assert not missing, f"declared in the header but not exported: {missing}"
    lib.pyi_abi_version.restype = C.c_int
    assert lib.pyi_abi_version() == 2


def test_ctypes_mirror_matches_header_layout():
    abi = importlib.import_module("pangeo-pyinterp_b200._abi")
    # pyi_axis_desc: 2 x i32, i64, 2 x i32, 5 x 8-byte unions, f64, i64, pointer
    assert C.sizeof(abi.AxisDesc) == 4 + 4 + 8 + 4 + 4 + 5 * 8 + 8 + 8 + 8
    assert C.sizeof(abi.GeometricConfig) == 7 * 4
    assert C.sizeof(abi.WindowedConfig) == 7 * 4
    assert C.sizeof(abi.WindowedNdConfig) == 10 * 4
    # value_mode is the LAST field of every configuration POD: the oracle's
    # checkers (orc_abi.h) read the same structs and ignore it
    for pod in (abi.GeometricConfig, abi.WindowedConfig, abi.WindowedNdConfig):
        assert pod._fields_[-1][0] == "value_mode"


def test_axes_are_host_objects(port):
    """Axis construction / description is host logic (no kernel): it must agree
    with the oracle's description of the same axis field by field."""
    pyi = importlib.import_module("pangeo-pyinterp_b200")
    lon = np.arange(1440) * 0.25
    a = pyi.Axis(lon, period=360.0)
    d = port.axis(lon, period=360.0).describe()
    assert len(a) == d["size"] == 1440
    assert a.is_regular() and a.is_periodic and a.is_ascending()
    assert a.min_value() == d["min_value"] and a.max_value() == d["max_value"]
    assert a.increment() == d["step"]
    irr = pyi.Axis(np.array([0.0, 1.0, 4.0, 8.0, 20.0, 50.0]))
    assert not irr.is_regular() and irr[2] == 4.0
    t = pyi.TemporalAxis(np.datetime64("2021-01-01", "ns") + np.arange(5) * np.timedelta64(1, "D"))
    assert t.dtype == np.dtype("datetime64[ns]") and len(t) == 5


def test_no_cpu_fallback():
    """Without a CUDA device the product path fails loudly (PYI_NO_DEVICE)."""
    pyi = importlib.import_module("pangeo-pyinterp_b200")
    if pyi.device_count() > 0:
        pytest.skip("a CUDA device is present")
    a = pyi.Axis(np.arange(10.0))
    with pytest.raises(RuntimeError, match="no CUDA device|CPU fallback"):
        a.find_indexes(np.array([1.5]))
    with pytest.raises(RuntimeError, match="no CUDA device|CPU fallback"):
        pyi.Grid(pyi.Axis(np.arange(4.0)), pyi.Axis(np.arange(3.0)), np.zeros((4, 3)))


def test_product_never_imports_the_oracle():
    """The oracle is test infrastructure: nothing under the package may import,
    load or execute it."""
    pkg = os.path.join(ROOT, "pangeo-pyinterp_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".hpp", ".h")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert "liboracle" not in text and "import orc" not in text and "oracle/" not in text, f


@pytest.mark.parametrize("model", ["linear", "c_spline", "c_spline_not_a_knot", "c_spline_periodic", "polynomial"])
def test_model_stencil_tables_against_the_oracle(model):
    """pyi_model_stencil (host arithmetic): the window weights the 6 x 6 tolerance kernels use for the
    fitting models that are linear in the node values, against the oracle's univariate model
    (univariate/{linear,cspline,cspline_not_a_knot,cspline_periodic,polynomial}.hpp) on six uniform
    nodes -- several steps and offsets (the polynomial table depends on the step: the reference's
    divided-difference update is aliased), random node values, t across the centre cell."""
    import orc
    o = orc.Oracle("reference" if orc.available("reference") else "port")
    abi = importlib.import_module("pangeo-pyinterp_b200._abi")
    lib = abi.load()
    rng = np.random.default_rng(4)
    for step in (0.25, 1.0, 2.5, 1.0 / 3.0, 30.0):
        table = np.zeros(36)
        st = lib.pyi_model_stencil(abi.SPATIAL_METHODS[model], step, table.ctypes.data_as(C.c_void_p))
        assert st == 0
        table = table.reshape(6, 6)
        for start in (0.0, -7.0 * step, 100.0 * step):
            xa = start + np.arange(6) * step
            ya = rng.normal(size=6) * 10.0
            for t in (0.03, 0.25, 0.5, 0.8, 0.97):
                want = o.univariate(model, xa, ya, xa[2] + t * step)
                got = float(sum(np.polyval(table[j][::-1], t) * ya[j] for j in range(6)))
                assert abs(got - want) <= 1e-12 * max(1.0, np.abs(ya).max()), (model, step, start, t, got, want)
    bad = np.zeros(36)
    assert lib.pyi_model_stencil(abi.SPATIAL_METHODS["akima"], 1.0, bad.ctypes.data_as(C.c_void_p)) != 0
