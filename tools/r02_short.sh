#!/bin/bash
cd /root/repo; mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q -k "decisions or every_cv2 or stage_map or inexact or golden or two_cascades or properties or tilted or randomized or queue_takes or edge or pipeline or submit" 2>&1 | tail -3
echo "short tiles (default):"; python tools/r02_lat.py 2>&1 | tail -2
echo "32-row tiles always:"; UGD_SHORT_TILES_BELOW=0 python tools/r02_lat.py 2>&1 | tail -2
timeout 300 python bench.py --steps 5 --warmup 3 --no-configs --no-cpu-baseline 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value'],1), round(d['e2e']['value'],1), d['parity']['mismatches'], d['single_frame_latency'])"
