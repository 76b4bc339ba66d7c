#!/bin/bash
# multi-GPU call: $1 = number of GPUs
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/m${N}_topo.txt 2>&1; nproc >> gpurun_out/m${N}_topo.txt; free -g >> gpurun_out/m${N}_topo.txt
timeout 600 ./tools/pcie_probe 5e7 > gpurun_out/m${N}_pcie_probe.txt 2>&1; echo "probe rc=$?"; cat gpurun_out/m${N}_pcie_probe.txt
timeout 300 python tools/e2e_small.py > gpurun_out/m${N}_e2e_small.txt 2>&1; cat gpurun_out/m${N}_e2e_small.txt
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q -s > gpurun_out/m${N}_multi_tests.log 2>&1; echo "multi tests rc=$?"; tail -8 gpurun_out/m${N}_multi_tests.log
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/m${N}_bench.json 2> gpurun_out/m${N}_bench.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/m${N}_bench.json; tail -5 gpurun_out/m${N}_bench.err
