#!/bin/bash
# all BASELINE configs once (scaled where large): usage tools/gpu_configs.sh <tag> "cfg scale" ...
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-cfg}; shift
for cs in "$@"; do
  set -- $cs
  SECONDS=0; timeout 900 python bench.py --config $1 --scale $2 --steps 2 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_$1.json 2> $OUT/${TAG}_$1.err
  echo "$1 scale=$2 rc=$?"; echo "  wall ${SECONDS}s"
  python - <<PY
import json
try:
    d=json.load(open("$OUT/${TAG}_$1.json"))
    print("  value %.4g  e2e %.4g  ms/step %.1f  frac %.3f  kernel_ms %.2f share %.3f launches %d nnz %d" % (d["value"], d["e2e"]["value"], d["ms_per_step"], d["roofline"]["frac"], d["roofline"]["kernel_ms_per_launch"], d["roofline"]["kernel_share_of_step"], d["gpu_launches"], d["config"]["nnz_train_per_gpu"]))
except Exception as e:
    print("  failed", e); print(open("$OUT/${TAG}_$1.err").read()[-1500:])
PY
done
