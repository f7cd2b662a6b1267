#!/bin/bash
cd /root/repo; mkdir -p gpurun_out
run() { echo -n "$1: "; env $2 timeout 300 python bench.py --steps 5 --warmup 3 --no-configs --no-cpu-baseline $3 2>gpurun_out/bgr_err.log | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(round(d['value'],1), 'fps  e2e', round(d['e2e']['value'],1), ' parity', d['parity']['mismatches'], {k:round(v['ms_per_launch'],4) for k,v in d['kernels'].items() if k in ('k_preprocess','k_area','k_equalize')})"; tail -2 gpurun_out/bgr_err.log; }
L8=/root/repo/ugproj-detector_b200/build/libugdet_it8.so
run "fused bytes it32" "A=1" "--cascade alt2 --scan-size 640x360"
run "fused atomic it32" "UGD_PRE_ATOMIC=1" "--cascade alt2 --scan-size 640x360"
run "fused bytes it8" "UGD_LIB=$L8" "--cascade alt2 --scan-size 640x360"
run "fused atomic it8" "UGD_LIB=$L8 UGD_PRE_ATOMIC=1" "--cascade alt2 --scan-size 640x360"
run "unfused atomic it8" "UGD_LIB=$L8 UGD_PRE_ATOMIC=1 UGD_NO_AREA_FUSE=1" "--cascade alt2 --scan-size 640x360"
run "unfused bytes it32" "UGD_NO_AREA_FUSE=1" "--cascade alt2 --scan-size 640x360"
