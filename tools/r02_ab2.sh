#!/bin/bash
# A/B of library variants (UGD_LIB) on the headline workload + decision tests with the default build
python -m pytest tests -m gpu -x -q -k "decisions or stage_map_720p or every_cv2" > gpurun_out/ab_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/ab_pytest.log
run() { echo -n "$1: "; env $2 python bench.py --steps 4 --warmup 3 --no-configs --no-cpu-baseline $3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(round(d['value'],1), 'fps  k_scan', round(d['kernels']['k_scan']['ms_per_launch'],3), 'ms  queued', d['counters_last_step']['queued'], 'raw', d['counters_last_step']['raw'])"; }
run "q4 D8" "A=1" ""
run "q3 D8" "UGD_LIB=$PWD/ugproj-detector_b200/build/libugdet_q3.so" ""
run "q2 D8" "UGD_LIB=$PWD/ugproj-detector_b200/build/libugdet_q2.so" ""
run "q4 D9" "A=1" "--dense-stages 9"
run "q4 D7" "A=1" "--dense-stages 7"
run "q4 D6" "A=1" "--dense-stages 6"
run "q3 D7" "UGD_LIB=$PWD/ugproj-detector_b200/build/libugdet_q3.so" "--dense-stages 7"
run "q4 D8 default cascade" "A=1" "--cascade default"
run "q4 D8 alt_tree" "A=1" "--cascade alt_tree"
