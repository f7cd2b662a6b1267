#!/bin/bash
# BGR front end: parity tests of the preprocess / scaled paths, then the shipped cfg's workload with and without the fused decimation
cd /root/repo; mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "preprocess or bgr or scaled or area or shim or golden or pipeline or submit" > gpurun_out/bgr_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/bgr_pytest.log
run() { echo -n "$1: "; env $2 timeout 300 python bench.py --steps 5 --warmup 3 --no-configs --no-cpu-baseline $3 2>gpurun_out/bgr_err.log | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(round(d['value'],1), 'fps  e2e', round(d['e2e']['value'],1), ' parity', d['parity']['mismatches'], {k:round(v['ms_per_launch'],4) for k,v in d['kernels'].items()})"; tail -2 gpurun_out/bgr_err.log; }
run "cfg mirror fused" "A=1" "--cascade alt2 --scan-size 640x360"
run "cfg mirror unfused" "UGD_NO_AREA_FUSE=1" "--cascade alt2 --scan-size 640x360"
run "alt gray 720p (rotated frames, 64 distinct)" "A=1" ""
