#!/bin/bash
# quick GPU visit: sweep-path parity test + a few bench settings ("table-mb sweep-chunk streams" triples)
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-q}; shift
timeout 900 python -m pytest tests -m gpu -x -q -k "sweep_table or golden or mma_path" > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $OUT/${TAG}_pytest.log
for cfg in "$@"; do
  set -- $cfg
  timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline --table-mb $1 --sweep-chunk $2 --sweep-streams $3 > $OUT/${TAG}_bench_$1_$2_$3.json 2> $OUT/${TAG}_bench_$1_$2_$3.err
  echo "table-mb=$1 chunk=$2 streams=$3 rc=$?"; python - <<PY
import json
try:
    d=json.load(open("$OUT/${TAG}_bench_$1_$2_$3.json"))
    print("  value %.4g  e2e %.4g  ms/step %.1f  frac %.3f  kernel_ms %.2f share %.3f launches %d" % (d["value"], d["e2e"]["value"], d["ms_per_step"], d["roofline"]["frac"], d["roofline"]["kernel_ms_per_launch"], d["roofline"]["kernel_share_of_step"], d["gpu_launches"]))
except Exception as e:
    print("  failed", e); print(open("$OUT/${TAG}_bench_$1_$2_$3.err").read()[-2000:])
PY
done
