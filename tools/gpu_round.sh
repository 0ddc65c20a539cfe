#!/bin/bash
# One GPU-box visit: parity tests, smoke, bench, ncu launch list + one full capture of the top kernel.
# Usage: tools/gpu_round.sh <tag> [bench args...]
TAG=${1:-r01}; shift
OUT=gpurun_out
mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $OUT/${TAG}_smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/${TAG}_pytest.log
tail -5 $OUT/${TAG}_pytest.log
timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' > $OUT/${TAG}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $OUT/${TAG}_smoke.log
timeout 900 python bench.py "$@" > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?"
cat $OUT/${TAG}_bench.json; tail -5 $OUT/${TAG}_bench.err
if [ -s $OUT/${TAG}_bench.json ]; then
  # launch list of ONE timed step at full size (skip data generation + 3 warm-up steps + the timed region's first launches)
  timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -s 900 -c 260 --csv --log-file $OUT/${TAG}_launches.csv \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline "$@" > $OUT/${TAG}_ncu_list.log 2>&1; echo "ncu list rc=$?"
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:accumulate_memo -s 40 -c 1 -o $OUT/${TAG}_accumulate -f \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline "$@" > $OUT/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
fi
