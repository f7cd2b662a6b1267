#!/bin/bash
cd /root/repo
run() { echo -n "$1: "; env $2 timeout 200 python bench.py --steps 5 --warmup 3 --no-configs --no-cpu-baseline $3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(round(d['value'],1), 'fps  e2e', round(d['e2e']['value'],1), ' k_scan', round(d['kernels']['k_scan']['ms_per_launch'],3), 'ms  parity', d['parity']['mismatches'], ' frac', round(d['roofline']['frac'],4))"; }
for t in 1200 1000000000; do
run "alt chunk_below=$t" "UGD_CHUNK_BELOW=$t" ""
run "alt2 chunk_below=$t" "UGD_CHUNK_BELOW=$t" "--cascade alt2"
run "default 1080p chunk_below=$t" "UGD_CHUNK_BELOW=$t" "--cascade default --size 1920x1080 --frames 64 --group 32"
run "4K chunk_below=$t" "UGD_CHUNK_BELOW=$t" "--size 3840x2160 --scale-factor 1.05 --min-size 24 --frames 16 --group 8"
run "cfg mirror chunk_below=$t" "UGD_CHUNK_BELOW=$t" "--cascade alt2 --scan-size 640x360"
done
