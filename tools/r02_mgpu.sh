#!/bin/bash
# multi-GPU check (gpurun --gpus N): the library-side NCCL gather against a 1-GPU run, then the bench at N ranks
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L | head -8
python -m pytest tests/test_gpu_multi.py -x -q -m gpu > gpurun_out/r02_pytest_mgpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r02_pytest_mgpu.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r02_bench_${N}gpu.json 2> gpurun_out/r02_bench_${N}gpu.err; echo "bench rc=$?"
tail -5 gpurun_out/r02_bench_${N}gpu.err
python - <<PY
import json
d=json.load(open('gpurun_out/r02_bench_${N}gpu.json'))
print('value',round(d['value'],1),'e2e',round(d['e2e']['value'],1),'ms/step',round(d['ms_per_step'],3),'parity',d['parity'])
for c in d.get('configs',[]): print(c['baseline_config'], round(c['value'],1), round(c['e2e']['value'],1), c['parity'])
PY
