#!/bin/bash
# the scan kernel compiled with -DUGD_CHECKED=1 (device-side bounds checks on every tile / SQ-plane read, trap on a miss)
# under the decision-level parity tests: all sizes, all cascade shapes, TMA and fallback staging
cd /root/repo; mkdir -p gpurun_out
L=/root/repo/ugproj-detector_b200/build/libugdet_checked.so
ls -la $L
UGD_LIB=$L timeout 1500 python -m pytest tests -m gpu -x -q --deselect tests/test_shim.py > gpurun_out/r02_checked_build_pytest.log 2>&1; echo "checked build pytest rc=$?"
tail -3 gpurun_out/r02_checked_build_pytest.log
UGD_LIB=$L UGD_NO_TMA=1 timeout 900 python -m pytest tests -m gpu -x -q -k "decisions or stage_map_720p or tilted" > gpurun_out/r02_checked_build_notma_pytest.log 2>&1; echo "checked build, fallback staging: pytest rc=$?"
tail -3 gpurun_out/r02_checked_build_notma_pytest.log

