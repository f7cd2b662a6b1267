#!/bin/bash
# A/B of the 32-row scan tiles (UGD_SCAN_R16=1 keeps 16 rows): decision-level parity, then the bench per cascade
cd /root/repo; mkdir -p gpurun_out
run() { echo -n "$1: "; env $2 timeout 200 python bench.py --steps 5 --warmup 3 --no-configs --no-cpu-baseline $3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(round(d['value'],1), 'fps  e2e', round(d['e2e']['value'],1), ' k_scan', round(d['kernels']['k_scan']['ms_per_launch'],3), 'ms  parity', d['parity']['mismatches'], ' frac', round(d['roofline']['frac'],4))"; }
timeout 900 python -m pytest tests -m gpu -x -q -k "decisions or every_cv2 or stage_map or inexact or golden or two_cascades or properties or tilted" > gpurun_out/tall_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/tall_pytest.log
run "alt tall" "A=1" ""
run "alt r16 " "UGD_SCAN_R16=1" ""
run "alt2 tall" "A=1" "--cascade alt2"
run "alt2 r16 " "UGD_SCAN_R16=1" "--cascade alt2"
run "alt_tree tall" "A=1" "--cascade alt_tree"
run "default 1080p tall" "A=1" "--cascade default --size 1920x1080 --frames 64 --group 32"
run "default 1080p r16 " "UGD_SCAN_R16=1" "--cascade default --size 1920x1080 --frames 64 --group 32"
