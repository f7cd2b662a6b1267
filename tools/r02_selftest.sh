#!/bin/bash
cd /root/repo; mkdir -p gpurun_out
L=/root/repo/ugproj-detector_b200/build/libugdet_checked.so
UGD_LIB=$L timeout 300 python -m pytest tests -m gpu -x -q -k "decisions" 2>&1 | tail -2
sed -n '/^# self-test of the checked build/,$p' tools/r02_checked.sh > /tmp/selftest.sh; L=$L bash /tmp/selftest.sh 2>&1 | tee gpurun_out/r02_checked_build_selftest.log
timeout 300 python bench.py --steps 5 --warmup 3 --no-configs > gpurun_out/lat_bench.json 2>gpurun_out/lat_err.log; echo "bench rc=$?"; tail -2 gpurun_out/lat_err.log
python -c "
import json; d=json.load(open('gpurun_out/lat_bench.json')); print(d['value'], d['e2e']['value'], d['single_frame_latency'])"
timeout 300 python bench.py --steps 5 --warmup 3 --no-configs --no-cpu-baseline --cascade alt2 --scan-size 640x360 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], d['e2e']['value'], d['single_frame_latency'])"
