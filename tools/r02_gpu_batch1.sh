#!/bin/bash
# Round-2 GPU batch 1 (one B200): GPU parity tests, bench both arms, launch list, ncu --set full of the shipping kernels.
# usage: gpurun --timeout 2400 -- 'bash tools/r02_gpu_batch1.sh'
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/r02_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r02_pytest_gpu.log
tail -5 gpurun_out/r02_pytest_gpu.log
python bench.py --steps 5 --warmup 3 > gpurun_out/r02_bench_1gpu.json 2> gpurun_out/r02_bench_1gpu.err; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_bench_reference_arm.json 2>/dev/null; echo "ref rc=$?"
B="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-configs"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r02_launches.csv $B > /dev/null 2>&1; echo "launch list rc=$?"
ncu --set full --clock-control none --import-source on -k regex:'k_resize|k_colsum|k_carry|k_integral' --launch-skip 12 -c 4 -o gpurun_out/r02_image_kernels -f $B > /dev/null 2>&1; echo "ncu image rc=$?"
ncu --set full --clock-control none --import-source on -k regex:k_scan --launch-skip 6 -c 2 -o gpurun_out/r02_k_scan -f $B > /dev/null 2>&1; echo "ncu scan rc=$?"
head -c 1500 gpurun_out/r02_bench_1gpu.json
