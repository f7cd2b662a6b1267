#!/bin/bash
run() { echo -n "$1: "; env $2 timeout 200 python bench.py --steps 5 --warmup 3 --no-configs --no-cpu-baseline $3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(round(d['value'],1), 'fps  e2e', round(d['e2e']['value'],1), ' pageable', round(d['e2e_pageable']['value'],1), ' k_scan', round(d['kernels']['k_scan']['ms_per_launch'],3), 'ms')"; }
timeout 600 python -m pytest tests -m gpu -x -q -k "decisions or stage_map_720p or two_cascades or pipeline or submit or batch or scaled or properties or golden" 2>&1 | tail -2
run "two scan streams" "A=1" ""
run "no group overlap" "UGD_NO_OVERLAP=1" ""
run "alt2" "A=1" "--cascade alt2"
run "1080p default" "A=1" "--cascade default --size 1920x1080 --frames 64 --group 32"
run "cfg mirror bgr scaled" "A=1" "--cascade alt2 --scan-size 640x360"
