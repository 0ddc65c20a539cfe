#!/bin/bash
# weak-scaling sweep on one box: usage [EXTRA='--config c5 --scale 0.125'] tools/gpu_scale.sh <tag> N1 N2 ...
OUT=gpurun_out; mkdir -p $OUT; TAG=$1; shift
for N in "$@"; do
  if [ "$N" = "1" ]; then
    timeout 900 python bench.py --gpus 1 --steps 5 --warmup 3 --no-cpu-baseline $EXTRA > $OUT/${TAG}_n$N.json 2> $OUT/${TAG}_n$N.err
  else
    timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N bench.py --gpus $N --steps 5 --warmup 3 $EXTRA > $OUT/${TAG}_n$N.json 2> $OUT/${TAG}_n$N.err
  fi
  echo "N=$N rc=$?"
  python - <<PY
import json
try:
    d=json.loads([l for l in open("$OUT/${TAG}_n$N.json") if l.startswith("{")][-1])
    print("  value %.4g e2e %.4g ms/step %.2f" % (d["value"], d["e2e"]["value"], d["ms_per_step"]))
except Exception as e:
    print("  failed", e); print(open("$OUT/${TAG}_n$N.err").read()[-1500:])
PY
done
