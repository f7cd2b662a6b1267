#!/bin/bash
# on the GPU box: bench summary for each library variant given (paths), e.g. tools/ab.sh ugproj-detector_b200/ab_*.so
# extra bench arguments through AB_ARGS, e.g. AB_ARGS="--cascade alt2"
for so in "$@"; do
  echo "== $so $AB_ARGS"
  UGD_LIB=$PWD/$so timeout 300 python bench.py --no-cpu-baseline --steps 3 $AB_ARGS 2>/dev/null | python -c "
import json,sys;d=json.loads(sys.stdin.read());print(round(d['value']),round(d['e2e']['value']),{k:round(v['ms_per_launch'],2) for k,v in d['kernels'].items()}, d['counters_last_step']['queued'], d['counters_last_step']['raw'])"
done
