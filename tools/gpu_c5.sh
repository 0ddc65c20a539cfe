#!/bin/bash
# C5-family runs: "config scale" pairs
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-c5}; shift
for cs in "$@"; do
  set -- $cs
  S0=$SECONDS
  timeout 900 python bench.py --config $1 --scale $2 --steps 3 --warmup 3 --no-cpu-baseline ${3:+--table-mb $3} > $OUT/${TAG}_$1_$2_$3.json 2> $OUT/${TAG}_$1_$2_$3.err
  echo "config=$1 scale=$2 table-mb=$3 rc=$? wall=$((SECONDS-S0))s"; python - <<PY
import json
try:
    d=json.load(open("$OUT/${TAG}_$1_$2_$3.json"))
    print("  value %.4g  e2e %.4g  ms/step %.1f  frac %.3f  kernel_ms %.2f share %.3f launches %d path %s" % (d["value"], d["e2e"]["value"], d["ms_per_step"], d["roofline"]["frac"], d["roofline"]["kernel_ms_per_launch"], d["roofline"]["kernel_share_of_step"], d["gpu_launches"], d["config"]["path"]))
except Exception as e:
    print("  failed", e); print(open("$OUT/${TAG}_$1_$2_$3.err").read()[-2000:])
PY
done
