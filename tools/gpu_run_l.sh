#!/bin/bash
mkdir -p gpurun_out
for w in trivariate quadrivariate; do
for v in 3 2 4 5; do
  export PYI_ND_MINB=$v
  timeout 600 python bench.py --no-cpu --no-secondary --steps 5 --warmup 3 --workload $w --points 1e8 > gpurun_out/l_${w}_$v.json 2> gpurun_out/l_${w}_$v.err
  python -c "
import json
d=json.loads([l for l in open('gpurun_out/l_${w}_$v.json') if l.startswith('{')][-1])
print('$w minb $v', round(d['value']), 'Mpts/s', round(d['roofline']['frac'],4))"
done; done
