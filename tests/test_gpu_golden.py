"""The CUDA path (public API -> C-ABI -> sm_100a kernels) against the committed
golden fixtures: the reference's own known-answer tests and the outputs of the
reference's headers on seeded cases (tests/golden/).  Nothing here reads
/root/reference or runs the CPU oracle.

Tolerances (BASELINE.json north_star): indexes bit-exact; float64 values
1e-12, float32 1e-5 relative with the scale guard max(|want|, 1)."""
import json
import os

import numpy as np
import pytest

import golden_util as gu

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
KA = json.load(open(os.path.join(HERE, "golden", "reference_known_answers.json")))


@pytest.mark.parametrize("name", gu.VECTOR_CASES)
def test_reference_vectors_on_gpu(pyi, name):
    d = gu.load_vectors(name)
    before = pyi.launch_count()
    got = gu.run_case_on_gpu(pyi, name, d)
    assert pyi.launch_count() > before, "no CUDA kernel was launched"
    gu.compare(name, got, d, exact=False)


def test_known_answer_lookups_on_gpu(pyi):
    for e in KA["axis_find_indexes"]:
        a = pyi.Axis(gu.axis_values(e), period=e["period"])
        assert a.find_indexes(np.array(e["coords"])).tolist() == e["pairs"], e["src"]
    for e in KA["axis_find_indexes_range"]:
        a = pyi.Axis(gu.axis_values(e), period=e["period"])
        got = a.find_indexes(gu.arange(e["coords_arange"]))
        assert np.array_equal(got[:, 0], gu.arange(e["i0_arange"]).astype(np.int64))
        assert np.array_equal(got[:, 1], gu.arange(e["i1_arange"]).astype(np.int64))
    for e in KA["axis_find_index"]:
        a = pyi.Axis(gu.axis_values(e), period=e["period"])
        for c, bounded, want in e["queries"]:
            assert a.find_index(np.array([c]), bounded=bounded)[0] == want, (e["src"], c)
    for e in KA["axis_find_index_i64"]:
        a = pyi.TemporalAxis(gu.axis_values_i64(e).astype("datetime64[s]"))
        if e["flip"]:
            a = a.flip(inplace=False)
        for c, bounded, want in e["queries"]:
            got = a.find_index(np.array([c], dtype=np.int64).astype("datetime64[s]"), bounded=bounded)
            assert got[0] == want, (e["src"], c)
    for e in KA["axis_find_indexes_i64"]:
        a = pyi.TemporalAxis(gu.axis_values_i64(e).astype("datetime64[ns]"))
        got = a.find_indexes(np.array(e["coords"], dtype=np.int64).astype("datetime64[ns]"))
        assert got.tolist() == e["pairs"], e["src"]


def _window_grid(pyi, e):
    xa, ya = np.array(e["xa"]), np.array(e["ya"])
    return pyi.Grid(pyi.Axis(xa), pyi.Axis(ya), np.array(e["za"], dtype=np.float64))


def test_known_answer_windowed_on_gpu(pyi):
    """GSL bicubic / bilinear vectors (bicubic.cpp:14-88, bilinear.cpp:14-65):
    the whole table is one window (shrink mode keeps every node)."""
    for e in KA["window_eval"]:
        grid = _window_grid(pyi, e)
        half_x, half_y = len(e["xa"]) - 1, len(e["ya"]) - 1
        cfg = pyi.windowed.Bivariate.bicubic() if e["method"] == "bicubic" else pyi.windowed.Bivariate.bilinear()
        cfg = cfg.with_half_window_size_x(half_x).with_half_window_size_y(half_y)
        cfg = cfg.with_boundary_mode(pyi.config.BoundaryConfig.shrink())
        got = pyi.core.bivariate(grid, np.array(e["xp"]), np.array(e["yp"]), cfg)
        assert np.all(np.abs(got - np.array(e["zp"])) <= e["tol"]), (e["src"], got)


def test_known_answer_geometric_cells_on_gpu(pyi):
    """biavriate.cpp cells as 2x2 grids.  The reference's tests use a Cartesian
    point; bilinear does not depend on the point type, so those vectors apply
    to the product path as they are.  idw / nearest on the product path are
    spherical (parity unpinned): only the corner / tie / equal
    value cases, which hold for any metric, are checked."""
    for e in KA["eval2d_cartesian"]:
        (x0, y0), (x1, y1) = e["p0"], e["p1"]
        q00, q01, q10, q11 = e["q"]
        xs, ys = np.array([x0, x1]), np.array([y0, y1])
        g = np.array([[q00, q01], [q10, q11]], dtype=np.float64)
        ax, ay = pyi.Axis(xs), pyi.Axis(ys)
        grid = pyi.Grid(ax, ay, g)
        cfg = getattr(pyi.geometric.Bivariate, e["method"])
        cfg = cfg(e["exp"]) if e["method"] == "idw" else cfg()
        for x, y, want in e["cases"]:
            on_corner = x in (x0, x1) and y in (y0, y1)
            if e["method"] != "bilinear" and not (on_corner or len(set(e["q"])) == 1):
                continue
            got = pyi.core.bivariate(grid, np.array([x]), np.array([y]), cfg)[0]
            assert abs(got - want) <= max(4 * np.spacing(abs(want)), 1e-12 * abs(want)), (e["src"], x, y, got)
