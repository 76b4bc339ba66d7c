"""Pins both CPU oracles (the plain-C restatement and, where built, the
reference's own headers compiled in place) against

  * tests/golden/reference_known_answers.json -- every known-answer vector the
    reference's own C++ / Python tests hold for this path (file:line in each
    entry), and
  * tests/golden/ref_vectors_*.npz -- outputs of the reference headers
    (oracle/_ref) on seeded inputs, generated in the build container by
    tests/golden/make_golden.py (committed next to the fixtures).

No GPU needed."""
import json
import os

import numpy as np
import pytest

import golden_util as gu

HERE = os.path.dirname(os.path.abspath(__file__))
KA = json.load(open(os.path.join(HERE, "golden", "reference_known_answers.json")))


@pytest.fixture(scope="module", params=["port", "reference"])
def orc_any(request):
    import orc
    if not orc.available(request.param):
        pytest.skip(f"{request.param} oracle not built")
    return orc.Oracle(request.param)


def test_known_answer_find_indexes(orc_any):
    for e in KA["axis_find_indexes"]:
        a = orc_any.axis(gu.axis_values(e), period=e["period"])
        got = a.find_indexes(np.array(e["coords"], dtype=np.float64))
        assert got.tolist() == e["pairs"], e["src"]
    for e in KA["axis_find_indexes_range"]:
        a = orc_any.axis(gu.axis_values(e), period=e["period"])
        c = gu.arange(e["coords_arange"])
        got = a.find_indexes(c)
        assert np.array_equal(got[:, 0], gu.arange(e["i0_arange"]).astype(np.int64)), e["src"]
        assert np.array_equal(got[:, 1], gu.arange(e["i1_arange"]).astype(np.int64)), e["src"]


def test_known_answer_find_index(orc_any):
    for e in KA["axis_find_index"]:
        a = orc_any.axis(gu.axis_values(e), period=e["period"])
        for c, bounded, want in e["queries"]:
            assert a.find_index(np.array([c]), bounded)[0] == want, (e["src"], c)
    for e in KA["axis_find_index_i64"]:
        a = orc_any.axis(gu.axis_values_i64(e))
        if e["flip"]:
            a.flip()
        for c, bounded, want in e["queries"]:
            assert a.find_index(np.array([c], dtype=np.int64), bounded)[0] == want, (e["src"], c)
    for e in KA["axis_find_indexes_i64"]:
        a = orc_any.axis(gu.axis_values_i64(e))
        got = a.find_indexes(np.array(e["coords"], dtype=np.int64))
        assert got.tolist() == e["pairs"], e["src"]


def test_known_answer_windows(orc_any):
    for e in KA["axis_window"]:
        a = orc_any.axis(gu.axis_values(e), period=e["period"])
        idx, centre = a.find_window(e["coord"], e["half"], e["boundary"])
        assert idx == e["indexes"] and list(centre) == e["centre"], e["src"]


def test_known_answer_evaluators(orc_any):
    m = {"bilinear": 0, "idw": 1, "nearest": 2}
    L = orc_any.lib
    for e in KA["eval2d_cartesian"]:
        for x, y, want in e["cases"]:
            got = L.orc_eval2d_f64(m[e["method"]], e["exp"], 1, x, y, *e["p0"], *e["p1"], *e["q"])
            # EXPECT_DOUBLE_EQ = within 4 ulp
            assert abs(got - want) <= 4 * np.spacing(abs(want)) + 0.0, (e["src"], x, y, got)
    # IdwExponentInfluence (biavriate.cpp:137-146)
    r1 = L.orc_eval2d_f64(1, 1, 1, 0.05, 0.05, 0.0, 0.0, 1.0, 1.0, 100.0, 0.0, 0.0, 0.0)
    r8 = L.orc_eval2d_f64(1, 8, 1, 0.05, 0.05, 0.0, 0.0, 1.0, 1.0, 100.0, 0.0, 0.0, 0.0)
    assert r8 > r1 and abs(r8 - 100.0) < 1e-6


def test_known_answer_axis_interp(orc_any):
    L = orc_any.lib
    for e in KA["axis_interp"]:
        mm = 0 if e["method"] == "linear" else 1
        for x, want in e["cases"]:
            got = [L.orc_axis_interp_f64(mm, float(x), float(e["x0"]), float(e["x1"]), e["y0"], e["y1"])]
            if not e.get("float_only"):
                got.append(L.orc_axis_interp_i64(mm, int(x), e["x0"], e["x1"], e["y0"], e["y1"]))
                got.append(L.orc_axis_interp_i64_f32(mm, int(x), e["x0"], e["x1"], e["y0"], e["y1"]))
            for g in got:
                if want is None:
                    assert np.isnan(g), e["src"]
                else:
                    assert g == pytest.approx(want, rel=1e-6), (e["src"], x)


def test_known_answer_windowed_evaluators(orc_any):
    for e in KA["window_eval"]:
        mm = 0 if e["method"] == "bicubic" else 1
        za = np.array(e["za"], dtype=np.float64)
        for x, y, want in zip(e["xp"], e["yp"], e["zp"]):
            got = orc_any.window_eval(mm, e["xa"], e["ya"], za, x, y)
            assert abs(got - want) <= e["tol"], (e["src"], x, y, got, want)


@pytest.mark.parametrize("name", gu.VECTOR_CASES)
def test_reference_vectors(orc_any, name):
    """Seeded input/output vectors produced by the reference's own headers."""
    if gu.CASES[name].get("ref_only") and orc_any.kind != "reference":
        pytest.skip("windowed N-D models are pinned by the reference build only")
    d = gu.load_vectors(name)
    got = gu.run_case_on_oracle(orc_any, name, d)
    gu.compare(name, got, d, exact=True)


def test_temporal_repr_golden(orc_any):
    """TemporalAxis::coordinate_repr (math/temporal_axis.hpp:80-83) texts of the
    reference build, committed as tests/golden/ref_temporal_repr.json."""
    cases = json.load(open(os.path.join(HERE, "golden", "ref_temporal_repr.json")))["cases"]
    assert len(cases) > 500
    for c in cases:
        assert orc_any.temporal_repr(c["value"], c["unit"], bool(c["timedelta"])) == c["text"], c


def test_temporal_repr_product_matches_golden():
    """The product's formatter (pyi_format_temporal: host logic of the C-ABI,
    no kernel, no GPU needed) against the same golden texts, and against
    numpy's own ISO text where numpy prints the same form."""
    import importlib
    pyi = importlib.import_module("pangeo-pyinterp_b200")
    cases = json.load(open(os.path.join(HERE, "golden", "ref_temporal_repr.json")))["cases"]
    axes = {}
    for c in cases:
        kind = "timedelta64" if c["timedelta"] else "datetime64"
        key = (kind, c["unit"])
        if key not in axes:
            pts = np.arange(3).astype(f"{kind}[{c['unit']}]")
            axes[key] = pyi.TemporalAxis(pts)
        assert axes[key].coordinate_repr(c["value"]) == c["text"], c
    ax = axes[("datetime64", "ns")]
    v = int(np.datetime64("2021-03-04T05:06:07.123456789", "ns").astype(np.int64))
    assert ax.coordinate_repr(v) == "2021-03-04T05:06:07.123456789" == str(np.datetime64(v, "ns"))
