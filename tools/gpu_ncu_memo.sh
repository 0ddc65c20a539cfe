#!/bin/bash
# one full ncu capture (with source) of accumulate_memo_kernel at C2: usage tools/gpu_ncu_memo.sh <tag> [lib.so]
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-ncu}
if [ -n "$2" ]; then export SSVI_B200_LIB=$PWD/$2; fi
timeout 900 ncu --set full --clock-control none --import-source on -k regex:accumulate_memo -s 40 -c 1 -o $OUT/${TAG}_memo -f \
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e > $OUT/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la $OUT/${TAG}_memo.ncu-rep
