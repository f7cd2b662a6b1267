#!/bin/bash
run() { echo -n "$1: "; env $2 timeout 200 python bench.py --steps 5 --warmup 3 --no-configs --no-cpu-baseline $3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(round(d['value'],1), 'fps  e2e', round(d['e2e']['value'],1), ' k_scan', round(d['kernels']['k_scan']['ms_per_launch'],3), 'ms')"; }
timeout 300 python -m pytest tests -m gpu -x -q -k "decisions or stage_map_720p or two_cascades or pipeline" 2>&1 | tail -2
run "concurrent scan launches" "A=1" ""
run "serial scan launches" "UGD_SCAN_SERIAL=1" ""
run "concurrent alt2" "A=1" "--cascade alt2"
run "serial alt2" "UGD_SCAN_SERIAL=1" "--cascade alt2"
run "concurrent 1080p default" "A=1" "--cascade default --size 1920x1080 --frames 64 --group 32"
run "serial 1080p default" "UGD_SCAN_SERIAL=1" "--cascade default --size 1920x1080 --frames 64 --group 32"
