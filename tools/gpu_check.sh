#!/bin/bash
# on the GPU box: parity tests, then a short bench summary.  usage: tools/gpu_check.sh [bench args]
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
timeout 300 python bench.py --no-cpu-baseline --steps 3 "$@" 2>gpurun_out/bench.err | tee gpurun_out/bench_last.json | python -c "
import json,sys;d=json.loads(sys.stdin.read());print(round(d['value']),round(d['e2e']['value']),{k:round(v['ms_per_launch'],2) for k,v in d['kernels'].items()}, d['counters_last_step']['queued'])"
tail -c 300 gpurun_out/bench.err
