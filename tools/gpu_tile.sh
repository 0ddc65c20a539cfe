#!/bin/bash
# L2-tiled launch order: parity tests + C5 at a fraction of its nnz for several tile-block counts
OUT=gpurun_out; mkdir -p $OUT; TAG=${1:-tile}; SCALE=${2:-0.2}; shift; shift
timeout 1200 python -m pytest tests -m gpu -x -q -k "tiled or sweep or batched_rows or e2e" > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 $OUT/${TAG}_pytest.log
timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_c2.json 2> $OUT/${TAG}_c2.err; echo "c2 rc=$?"
python -c "
import json; d=json.load(open('$OUT/${TAG}_c2.json')); print('  c2 value %.4g e2e %.4g frac %.3f launches %d' % (d['value'], d['e2e']['value'], d['roofline']['frac'], d['gpu_launches']))"
for nb in "$@"; do
  S0=$SECONDS
  timeout 900 python bench.py --config c5 --scale $SCALE --steps 3 --warmup 3 --no-cpu-baseline --tile-blocks $nb > $OUT/${TAG}_c5_nb$nb.json 2> $OUT/${TAG}_c5_nb$nb.err
  echo "tile-blocks=$nb rc=$? wall=$((SECONDS-S0))s"; python - <<PY
import json
try:
    d=json.load(open("$OUT/${TAG}_c5_nb$nb.json"))
    print("  value %.4g  e2e %.4g  ms/step %.1f  frac %.3f  kernel_ms %.2f share %.3f launches %d path %s" % (d["value"], d["e2e"]["value"], d["ms_per_step"], d["roofline"]["frac"], d["roofline"]["kernel_ms_per_launch"], d["roofline"]["kernel_share_of_step"], d["gpu_launches"], d["config"]["path"]))
except Exception as e:
    print("  failed", e); print(open("$OUT/${TAG}_c5_nb$nb.err").read()[-1500:])
PY
done
