#!/bin/bash
# config 5 at FULL size on one GPU (1e9 observed entries, 8e8 train): device-resident arm only
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-c5full}; shift
free -g | head -2; nvidia-smi --query-gpu=memory.total,memory.used --format=csv,noheader
for nb in "$@"; do
  S0=$SECONDS
  timeout 1500 python bench.py --config c5 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e --tile-blocks $nb > $OUT/${TAG}_nb$nb.json 2> $OUT/${TAG}_nb$nb.err
  echo "tile-blocks=$nb rc=$? wall=$((SECONDS-S0))s"; python - <<PY
import json
try:
    d=json.load(open("$OUT/${TAG}_nb$nb.json"))
    print("  value %.4g  ms/step %.1f  frac %.3f  kernel_ms %.2f share %.3f launches %d path %s" % (d["value"], d["ms_per_step"], d["roofline"]["frac"], d["roofline"]["kernel_ms_per_launch"], d["roofline"]["kernel_share_of_step"], d["gpu_launches"], d["config"]["path"]))
except Exception as e:
    print("  failed", e); print(open("$OUT/${TAG}_nb$nb.err").read()[-1500:])
PY
done
