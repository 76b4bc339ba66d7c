#!/bin/bash
# dress rehearsal of the driver's round-end commands
mkdir -p gpurun_out
python __graft_entry__.py smoke > gpurun_out/k_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/k_smoke.log
N=${1:-2}
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --impl reference --gpus $N --steps 3 --warmup 1 > gpurun_out/k_ref_$N.json 2> gpurun_out/k_ref_$N.err; echo "ref arm rc=$?"; cut -c1-200 gpurun_out/k_ref_$N.json
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/k_bench_$N.json 2> gpurun_out/k_bench_$N.err ) 2>&1 | grep real; echo "bench rc=$?"
python - <<PY
import json
lines=[l for l in open('gpurun_out/k_bench_$N.json')]
print('stdout lines:', len(lines), 'non-json:', [l[:60] for l in lines if not l.startswith('{')])
d=json.loads([l for l in lines if l.startswith('{')][-1])
print('value',d['value'],'frac',d['roofline']['frac'],'parity',d['config']['parity'])
e=d['e2e']; print('e2e',e['value'],e['host_link_gbs'],e['frac_of_probe'],'live',e.get('live_probe_gbs'),e.get('frac_of_live_probe'),'pageable',e['pageable']['value'],'ranks',e['ranks'])
for k,v in d['secondary'].items():
    if 'error' in v: print(k, v['error']); continue
    print(k, round(v['value']), round(v['roofline']['frac'],4), v['parity']['bit_identical_fraction'], v['parity']['max_abs_err'], v['roofline']['traffic'] is not None)
PY
tail -3 gpurun_out/k_bench_$N.err
