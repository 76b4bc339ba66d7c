#!/usr/bin/env python
"""Generates tests/golden/ref_vectors_<case>.npz.

Run in the build container, where /root/reference exists and
`make -C oracle ref` has compiled the reference's OWN headers
(math/axis.hpp, geometric/bivariate.hpp, multivariate.hpp, bicubic.hpp,
cspline.hpp, cache.hpp ... see oracle/ref/ref_driver.cpp) into
oracle/_ref/liboracle_ref.so:

    python tests/golden/make_golden.py

Each fixture stores the OUTPUTS of the reference build for one seeded case of
tests/golden_util.py plus checksums of the inputs they were computed from (the
inputs themselves are rebuilt deterministically by golden_util.build_case).
The fixtures travel to the GPU box; /root/reference does not."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
for p in (ROOT, os.path.join(ROOT, "oracle"), os.path.dirname(HERE)):
    sys.path.insert(0, p)

import golden_util as gu  # noqa: E402
import orc  # noqa: E402


def main():
    if not orc.available("reference"):
        raise SystemExit("oracle/_ref/liboracle_ref.so is missing: run `make -C oracle ref` where /root/reference exists")
    o = orc.Oracle("reference")
    total = 0
    for name in gu.VECTOR_CASES:
        d = gu.build_case(name)
        got = gu.run_case_on_oracle(o, name, d)
        extra = {"sum_" + k: np.int64(gu.checksum(d[k])) for k in ("x", "y", "z", "u", "grid") if k in d}
        path = gu.fixture_path(name)
        np.savez_compressed(path, **got, **extra)
        total += os.path.getsize(path)
        first = next(iter(got.values()))
        nan = int(np.isnan(first).sum()) if first.dtype.kind == "f" else int((first < 0).sum())
        print(f"{name:32s} {os.path.getsize(path):8d} B  n={first.shape[0]}  undefined={nan}")
    print(f"total {total / 1024:.1f} KiB")
    temporal_repr_fixture(o)


def temporal_repr_cases():
    """(unit, is_timedelta, value) triples: epoch neighbourhood, negative values
    (floor semantics), leap days, far dates, sub-second fractions."""
    rng = np.random.Generator(np.random.PCG64(7))
    per_day = {"D": 1, "h": 24, "m": 1440, "s": 86400, "ms": 86400 * 10**3, "us": 86400 * 10**6,
               "ns": 86400 * 10**9}
    cases = []
    for unit in ("Y", "M", "W", "D", "h", "m", "s", "ms", "us", "ns", "ps", "fs", "as"):
        vals = [0, 1, -1, 59, -59, 60, 3599, -3601, 86399, -86401, 10**9 + 7, -(10**9) - 7]
        if unit in per_day:
            for day in (18628, 11016, -25567, 19051, 47482, -141427):  # 2021-01-01, 2000-02-29, 1900-01-01 ...
                vals += [day * per_day[unit], day * per_day[unit] - 1, day * per_day[unit] + per_day[unit] // 3]
        span = {"Y": 5000, "M": 60000, "W": 200000}.get(unit, 2**62)
        vals += [int(v) for v in rng.integers(-span, span, 24)]
        for v in vals:
            if unit in ("Y", "M", "W") and abs(v) > span:
                continue
            cases.append((unit, 0, int(v)))
        cases += [(unit, 1, int(v)) for v in (0, 1, -1, 12345, -(2**40))]
    return cases


def temporal_repr_fixture(o):
    import json
    out = [{"unit": u, "timedelta": td, "value": v, "text": o.temporal_repr(v, u, bool(td))}
           for u, td, v in temporal_repr_cases()]
    path = os.path.join(HERE, "ref_temporal_repr.json")
    with open(path, "w") as f:
        json.dump({"source": "oracle/_ref: dateutils::datetime64_to_string / timedelta64_to_string "
                             "(cxx/include/pyinterp/dateutils.hpp:729-812) = TemporalAxis::coordinate_repr",
                   "cases": out}, f, indent=0)
    print(f"ref_temporal_repr.json: {len(out)} cases, {os.path.getsize(path)} B")


if __name__ == "__main__":
    main()
