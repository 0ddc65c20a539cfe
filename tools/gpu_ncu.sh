#!/bin/bash
# ncu captures of the sweep kernels at full C2 scale. usage: tools/gpu_ncu.sh <tag> <kernel-regex> [bench args]
OUT=gpurun_out; mkdir -p $OUT; TAG=$1; KR=$2; shift; shift
timeout 600 python bench.py --steps 1 --warmup 3 --no-cpu-baseline "$@" > $OUT/${TAG}_plain.json 2> $OUT/${TAG}_plain.err && \
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:$KR -s 40 -c 1 -o $OUT/${TAG} -f \
   python bench.py --steps 1 --warmup 3 --no-cpu-baseline "$@" > $OUT/${TAG}_ncu.log 2>&1
echo "rc=$?"; tail -3 $OUT/${TAG}_ncu.log
