#!/bin/bash
# Round-2 evidence on one B200: every GPU test, both bench arms, the shipped cfg's workload, the launch list and
# ncu --set full captures of the shipping kernels (each ncu run only after the same command exited 0 without ncu).
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r02_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/r02_pytest_gpu.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r02_bench_1gpu.json 2> gpurun_out/r02_bench_1gpu.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/r02_bench_reference_arm.json 2>/dev/null; echo "ref rc=$?"
timeout 300 python bench.py --steps 5 --warmup 3 --cascade alt2 --scan-size 640x360 --no-configs > gpurun_out/r02_bench_cfg_mirror_scaled.json 2>/dev/null; echo "cfg mirror rc=$?"
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-configs"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches.csv $B > /dev/null 2>&1; echo "launch list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_resize|k_carry|k_integral' --launch-skip 9 -c 3 -o gpurun_out/r02_image_kernels -f $B > /dev/null 2>&1; echo "ncu image rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_scan --launch-skip 6 -c 2 -o gpurun_out/r02_k_scan -f $B > /dev/null 2>&1; echo "ncu scan rc=$?"
B2="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-configs --cascade alt2 --scan-size 640x360"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_preprocess|k_equalize' --launch-skip 6 -c 2 -o gpurun_out/r02_bgr_kernels -f $B2 > /dev/null 2>&1; echo "ncu bgr rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/r02_bench_1gpu.json'))
print('value',round(d['value'],1),'e2e',round(d['e2e']['value'],1),'pageable',round(d['e2e_pageable']['value'],1),'parity',d['parity']['mismatches'],'/',d['parity']['frames_checked'])
print('roofline frac',round(d['roofline']['frac'],4),'peak',round(d['roofline']['peak'],0))
print('stages', {k:(round(v['ms'],4), round(v['frac'],3)) for k,v in d['stages'].items()})
print({k:round(v['ms_per_launch'],4) for k,v in d['kernels'].items()})
for c in d.get('configs',[]): print(c['baseline_config'], round(c['value'],1), round(c['e2e']['value'],1), c['parity']['mismatches'], c['parity']['frames_checked'])
PY
