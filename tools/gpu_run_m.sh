#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/m_tests.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/m_tests.log
timeout 300 python tools/e2e_small.py > gpurun_out/m_e2e_small.txt 2>&1; cat gpurun_out/m_e2e_small.txt
python __graft_entry__.py smoke 2>&1 | tail -1
( time python bench.py > gpurun_out/m_bench.json 2> gpurun_out/m_bench.err ) 2>&1 | grep real; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/m_bench.json') if l.startswith('{')][-1])
print('value',d['value'],'frac',d['roofline']['frac'],'launches',d['gpu_launches'],'parity',d['config']['parity']['points'])
e=d['e2e']; print('e2e',e['value'],e['host_link_gbs'],e['frac_of_probe'],'pageable',e['pageable']['value'])
print('cpu',d['cpu_baseline']['value'])
for k,v in d['secondary'].items():
    if 'error' in v: print(k, v['error']); continue
    print(k, round(v['value']), round(v['roofline']['frac'],4), v['parity']['bit_identical_fraction'], v['parity']['max_abs_err'])
    for s in v.get('sweep',[]): print('    ', s['points_per_gpu'], round(s['value']), s['ms_per_step'])
PY
