#!/bin/bash
# parameter sweep of the scan with 32-row tiles: survivors below which a tile goes to the queue, in-tile stages
cd /root/repo; mkdir -p gpurun_out
run() { echo -n "$1: "; env $2 timeout 200 python bench.py --steps 5 --warmup 3 --no-configs --no-cpu-baseline $3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(round(d['value'],1), 'fps  k_scan', round(d['kernels']['k_scan']['ms_per_launch'],3), 'ms  parity', d['parity']['mismatches'])"; }
for tm in 16 32 48 64 96 128; do run "alt tile_min $tm" "UGD_TILE_MIN=$tm" ""; done
for D in 8 10 11; do run "alt D $D" "A=1" "--dense-stages $D"; done
for tm in 16 64 128; do run "alt2 tile_min $tm" "UGD_TILE_MIN=$tm" "--cascade alt2"; done
for D in 8 10 12; do run "alt2 D $D" "A=1" "--cascade alt2 --dense-stages $D"; done
for k in 1 3 4; do run "alt kmin $k" "UGD_SCAN_KMIN=$k" ""; done
