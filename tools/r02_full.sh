#!/bin/bash
# full single-GPU check: every GPU test, then the headline bench with all legs
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r02_pytest_gpu.log
python bench.py --steps 5 --warmup 3 > gpurun_out/r02_bench_1gpu.json 2> gpurun_out/r02_bench_1gpu.err; echo "bench rc=$?"; tail -3 gpurun_out/r02_bench_1gpu.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02_bench_1gpu.json'))
print('value',round(d['value'],1),'e2e',round(d['e2e']['value'],1),'pageable',round(d['e2e_pageable']['value'],1),'parity',d['parity']['mismatches'],'/',d['parity']['frames_checked'])
print('roofline frac',round(d['roofline']['frac'],4),'peak',round(d['roofline']['peak'],0))
print('stages', {k:(round(v['ms'],4), round(v['frac'],3)) for k,v in d['stages'].items()})
print({k:round(v['ms_per_launch'],4) for k,v in d['kernels'].items()})
for c in d.get('configs',[]): print(c['baseline_config'], round(c['value'],1), round(c['e2e']['value'],1), c['parity']['mismatches'], c['parity']['frames_checked'])
PY
