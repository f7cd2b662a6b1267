#!/bin/bash
# ncu --set full of the two k_scan launches of one launch group (after the program ran clean without ncu)
mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-configs"
$B > /dev/null 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:k_scan --launch-skip 6 -c 2 -o gpurun_out/${1:-scan} -f $B > /dev/null 2>&1; echo "ncu scan rc=$?"
