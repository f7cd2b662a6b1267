#!/bin/bash
cd /root/repo; mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q -k "pyramid or integral or level or decisions or stage_map_720p or golden or properties or no_alloc" 2>&1 | tail -3
run() { echo -n "$1: "; env $2 timeout 200 python bench.py --steps 5 --warmup 3 --no-configs --no-cpu-baseline $3 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read()); print(round(d['value'],1), 'fps  e2e', round(d['e2e']['value'],1), ' parity', d['parity']['mismatches'], ' frac', round(d['roofline']['frac'],4), {k:round(v['ms_per_launch'],4) for k,v in d['kernels'].items()}, {k:(round(v['ms'],4), round(v['frac'],3)) for k,v in d['stages'].items()})"; }
run "alt" "A=1" ""
run "alt2" "A=1" "--cascade alt2"
run "default 1080p" "A=1" "--cascade default --size 1920x1080 --frames 64 --group 32"
run "4K" "A=1" "--size 3840x2160 --scale-factor 1.05 --min-size 24 --frames 16 --group 8"
