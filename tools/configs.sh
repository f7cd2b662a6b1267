#!/bin/bash
# on the GPU box: the five BASELINE configs, one bench line each -> gpurun_out/configs.jsonl
out=gpurun_out/configs.jsonl; : > $out
run() { echo "== $*"; timeout 600 python bench.py --steps 3 "$@" 2>gpurun_out/cfg.err | tee -a $out | python -c "
import json,sys;d=json.loads(sys.stdin.read());print(round(d['value'],1),'fps e2e',round(d['e2e']['value'],1),'cpu',d['cpu_baseline'] and round(d['cpu_baseline']['value'],2),{k:round(v['ms_per_launch'],2) for k,v in d['kernels'].items()})" || tail -5 gpurun_out/cfg.err; }
run --cascade alt
run --cascade alt2
run --cascade default --size 1920x1080 --frames 128 --group 32
run --cascade alt_tree
run --cascade alt --size 3840x2160 --scale-factor 1.05 --min-size 24 --frames 16 --group 8
