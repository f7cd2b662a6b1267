#!/bin/bash
cd /root/repo; mkdir -p gpurun_out
nproc
run() { echo -n "$1: "; env $2 timeout 300 python bench.py --steps 5 --warmup 3 --no-configs --no-cpu-baseline $3 2>gpurun_out/page_err.log | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(round(d['value'],1), 'fps  e2e', round(d['e2e']['value'],1), ' pageable', round(d['e2e_pageable']['value'],1), ' parity', d['parity']['mismatches'])"; tail -2 gpurun_out/page_err.log; }
timeout 600 python -m pytest tests -m gpu -x -q -k "batch or pipeline or submit or scaled or shim or truncat" 2>&1 | tail -3
run "alt pin ring 4 threads" "A=1" ""
run "alt no ring" "UGD_NO_PIN_RING=1" ""
run "alt ring 12 threads" "UGD_COPY_THREADS=12" ""
run "cfg mirror ring 12 threads" "UGD_COPY_THREADS=12" "--cascade alt2 --scan-size 640x360"
run "cfg mirror ring" "A=1" "--cascade alt2 --scan-size 640x360"
run "cfg mirror no ring" "UGD_NO_PIN_RING=1" "--cascade alt2 --scan-size 640x360"
