#!/bin/bash
# usage: tools/gpu_ncu2.sh <tag> <kernel-regex> <skip> <bench args...>
OUT=gpurun_out; mkdir -p $OUT; TAG=$1; KR=$2; SK=$3; shift; shift; shift
timeout 900 python bench.py --steps 1 --warmup 3 --no-cpu-baseline "$@" > $OUT/${TAG}_plain.json 2> $OUT/${TAG}_plain.err && \
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:$KR -s $SK -c 1 -o $OUT/${TAG} -f \
   python bench.py --steps 1 --warmup 3 --no-cpu-baseline "$@" > $OUT/${TAG}_ncu.log 2>&1
echo "rc=$?"; python - <<PY
import json
d=json.load(open("$OUT/${TAG}_plain.json"))
print("value %.4g e2e %.4g ms/step %.1f frac %.3f" % (d["value"], d["e2e"]["value"], d["ms_per_step"], d["roofline"]["frac"]))
PY
tail -2 $OUT/${TAG}_ncu.log
