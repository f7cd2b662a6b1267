#!/bin/bash
# quick GPU check of a k_scan change: the decision-level parity tests, then the headline bench without the legs
# (every step under its own timeout: a kernel that hangs must not hold the box)
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "decisions or every_cv2 or stage_map or inexact or golden or two_cascades or properties or tilted" > gpurun_out/quick_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/quick_pytest.log
timeout 300 python bench.py --steps 5 --warmup 3 --no-configs > gpurun_out/quick_bench.json 2> gpurun_out/quick_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/quick_bench.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/quick_bench.json'))
print('value',round(d['value'],1),'e2e',round(d['e2e']['value'],1),'pageable',round(d['e2e_pageable']['value'],1),'parity',d['parity']['mismatches'],'/',d['parity']['frames_checked'])
print('roofline frac',round(d['roofline']['frac'],4),'peak',round(d['roofline']['peak'],0),'nominal frac',round(d['roofline']['frac_of_nominal'],4))
print({k:round(v['ms_per_launch'],4) for k,v in d['kernels'].items()})
PY
for c in alt2 default alt_tree; do timeout 200 python bench.py --steps 3 --warmup 3 --no-configs --no-cpu-baseline --cascade $c 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print('$c', round(d['value'],1), round(d['kernels']['k_scan']['ms_per_launch'],3))"; done
echo -n "no TMA: "; UGD_NO_TMA=1 timeout 200 python bench.py --steps 3 --warmup 3 --no-configs --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['kernels']['k_scan']['ms_per_launch'],3))"
