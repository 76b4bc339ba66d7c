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


if __name__ == "__main__":
    main()
