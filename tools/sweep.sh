#!/bin/bash
# on the GPU box: k_scan time for a grid of dense-stage counts and tile_min thresholds.  usage: tools/sweep.sh "6 7 8 9" "16 32 64" [bench args]
DS="$1"; TM="$2"; shift 2
for d in $DS; do for t in $TM; do
  echo -n "dense $d tile_min $t: "
  UGD_TILE_MIN=$t timeout 300 python bench.py --no-cpu-baseline --steps 3 --dense-stages $d "$@" 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read());print(round(d['value']),round(d['kernels']['k_scan']['ms_per_launch'],2), d['counters_last_step']['queued'])"
done; done
