#!/bin/bash
run() { echo -n "$1: "; env $2 python bench.py --steps 4 --warmup 3 --no-configs --no-cpu-baseline $3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(round(d['value'],1), 'fps  k_scan', round(d['kernels']['k_scan']['ms_per_launch'],3), 'ms  queued', d['counters_last_step']['queued'], 'raw', d['counters_last_step']['raw'])"; }
python -m pytest tests -m gpu -x -q -k "decisions or stage_map_720p or every_cv2 or inexact" > gpurun_out/ab_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/ab_pytest.log
for c in alt alt2 default alt_tree; do
run "$c D8" "A=1" "--cascade $c"
run "$c D9" "A=1" "--cascade $c --dense-stages 9"
run "$c D10" "A=1" "--cascade $c --dense-stages 10"
done
