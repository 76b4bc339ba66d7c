#!/bin/bash
# Times the bilinear fast kernel for each tuning variant (PYI_FAST_VARIANT).
for v in ${VARIANTS:-4n 4p 3n 3p 2n 2p}; do
  echo -n "$v "
  PYI_FAST_VARIANT=$v python bench.py --points ${POINTS:-1e9} --e2e-points 1e6 --steps ${STEPS:-20} --warmup 3 --no-cpu 2>&1 | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); print(round(d['value']), 'Mpts/s  min_ms', round(d['config']['kernel_ms_min'],3), 'med_ms', round(d['config']['kernel_ms_median'],3), d['clocks'])
"
done
