#!/bin/bash
# A/B builds of the kernels on one box: usage tools/gpu_ab.sh <tag> <config> <scale> lib1.so lib2.so ...
OUT=gpurun_out; mkdir -p $OUT; TAG=$1; CFG=$2; SCALE=$3; shift 3
for lib in "$@"; do
  name=$(basename $lib .so)
  SECONDS=0
  SSVI_B200_LIB=$PWD/$lib timeout 600 python bench.py --config $CFG --scale $SCALE --steps 3 --warmup 3 --no-cpu-baseline --no-e2e \
    > $OUT/${TAG}_${name}.json 2> $OUT/${TAG}_${name}.err
  echo "$name rc=$? wall ${SECONDS}s"
  python - <<PY
import json
try:
    d=json.load(open("$OUT/${TAG}_${name}.json"))
    print("  value %.4g  ms/step %.2f  frac %.3f  kernel_ms %.3f share %.3f" % (d["value"], d["ms_per_step"], d["roofline"]["frac"], d["roofline"]["kernel_ms_per_launch"], d["roofline"]["kernel_share_of_step"]))
except Exception as e:
    print("  failed", e); print(open("$OUT/${TAG}_${name}.err").read()[-1200:])
PY
done
