#!/bin/bash
# A/B of k_scan switches on the headline workload: prints fps and k_scan ms per launch group for each setting
run() { echo -n "$1: "; env $2 python bench.py --steps 4 --warmup 3 --no-configs --no-cpu-baseline $3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(round(d['value'],1), 'fps  k_scan', round(d['kernels']['k_scan']['ms_per_launch'],3), 'ms  queued', d['counters_last_step']['queued'])"; }
run "chunks D8" "A=1" ""
run "old queue D8" "UGD_NO_QCHUNKS=1" ""
run "old queue D10" "UGD_NO_QCHUNKS=1" "--dense-stages 10"
run "old queue D12" "UGD_NO_QCHUNKS=1" "--dense-stages 12"
run "old queue D14" "UGD_NO_QCHUNKS=1" "--dense-stages 14"
run "chunks D10" "A=1" "--dense-stages 10"
run "chunks D6" "A=1" "--dense-stages 6"
run "old queue D8 tilemin 16" "UGD_NO_QCHUNKS=1 UGD_TILE_MIN=16" ""
run "old queue D8 tilemin 64" "UGD_NO_QCHUNKS=1 UGD_TILE_MIN=64" ""
run "old queue D12 tilemin 8" "UGD_NO_QCHUNKS=1 UGD_TILE_MIN=8" "--dense-stages 12"
