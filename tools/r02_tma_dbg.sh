#!/bin/bash
for m in 0 1 2 3; do echo "== UGD_TMA_STEPS=$m"; UGD_TMA_STEPS=$m timeout 120 python -m pytest tests -m gpu -x -q -k "decisions and alt2" 2>&1 | tail -3; done
