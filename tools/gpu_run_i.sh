#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/i_tests.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/i_tests.log
timeout 900 python bench.py --no-cpu > gpurun_out/i_bench.json 2> gpurun_out/i_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/i_bench.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/i_bench.json') if l.startswith('{')][-1])
print('value',d['value'],'frac',d['roofline']['frac'],'e2e',d['e2e']['value'])
for k,v in d['secondary'].items():
    if 'error' in v: print(k, v['error']); continue
    print(k, round(v['value']), round(v['roofline']['frac'],4), v['parity']['bit_identical_fraction'], v['parity']['max_abs_err'])
PY
