#!/bin/bash
# GPU visit: subsampled-sweep planner test + bench lines for reference-style minibatch sweeps (B=128) and full sweeps
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-sub}
timeout 900 python -m pytest tests -m gpu -x -q -k "subsampled or batched_rows or sweep_table" > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $OUT/${TAG}_pytest.log
for bs in 128 0; do
  timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --batch-size $bs > $OUT/${TAG}_bench_b$bs.json 2> $OUT/${TAG}_bench_b$bs.err
  echo "batch-size=$bs rc=$?"; python - <<PY
import json
try:
    d=json.load(open("$OUT/${TAG}_bench_b$bs.json"))
    print("  value %.4g  e2e %.4g  ms/step %.2f  frac %.3f  kernel_ms %.3f share %.3f launches %d units/step %d" % (d["value"], d["e2e"]["value"], d["ms_per_step"], d["roofline"]["frac"], d["roofline"]["kernel_ms_per_launch"], d["roofline"]["kernel_share_of_step"], d["gpu_launches"], d["config"]["units_per_step_per_gpu"]))
except Exception as e:
    print("  failed", e); print(open("$OUT/${TAG}_bench_b$bs.err").read()[-2000:])
PY
done
