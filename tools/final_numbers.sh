#!/bin/bash
# on the GPU box: every bench line the docs quote -> gpurun_out/ (copy the ones to keep into profiles/)
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py > gpurun_out/bench_1gpu.json 2> gpurun_out/bench_1gpu.err; tail -c 400 gpurun_out/bench_1gpu.err
python bench.py --impl reference > gpurun_out/bench_reference_arm.json 2>/dev/null
bash tools/configs.sh
python bench.py --steps 3 --cascade alt2 --size 1280x720 --scan-size 640x360 > gpurun_out/bench_cfg_mirror.json 2>/dev/null
: > gpurun_out/other_cascades.jsonl
for c in smile upperbody eye_tree_eyeglasses frontalcatface_extended eye; do
  timeout 600 python bench.py --steps 3 --cascade $c 2>/dev/null >> gpurun_out/other_cascades.jsonl
done
python - <<'PY'
import json
for f in ("bench_1gpu.json", "bench_reference_arm.json", "bench_cfg_mirror.json"):
    d = json.loads(open("gpurun_out/" + f).read()); print(f, round(d["value"], 1), d.get("e2e", {}).get("value"), (d.get("cpu_baseline") or {}).get("value"))
for l in open("gpurun_out/other_cascades.jsonl"):
    d = json.loads(l); print(d["config"]["workload"][:40], round(d["value"]), round(d["e2e"]["value"]), d["cpu_baseline"] and round(d["cpu_baseline"]["value"], 1))
PY
