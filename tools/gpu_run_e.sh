#!/bin/bash
# bicubic K4s: parity tests, then throughput against K4w and over resident blocks per SM
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "bicubic or windowed or golden" > gpurun_out/e_tests.log 2>&1; echo "tests rc=$?"; tail -6 gpurun_out/e_tests.log
B="python bench.py --no-cpu --no-secondary --steps 5 --warmup 3 --workload bicubic --points 1e8"
for v in 6; do
  if [ $v = w ]; then export PYINTERP_B200_BICUBIC=weights; unset PYINTERP_B200_S_MINB; else unset PYINTERP_B200_BICUBIC; export PYINTERP_B200_S_MINB=$v; fi
  timeout 300 $B > gpurun_out/e_bicubic_$v.json 2> gpurun_out/e_bicubic_$v.err
  python -c "
import json,sys
d=json.loads([l for l in open('gpurun_out/e_bicubic_$v.json') if l.startswith('{')][-1])
print('variant $v', round(d['value']), 'Mpts/s', d['ms_per_step'], 'ms', 'e2e', round(d['e2e']['value']))"
done
