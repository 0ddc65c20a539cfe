#!/bin/bash
# config 4 (1e8 observed entries in total, 4-way, diagonal, SSVI_TF_simple), strong scaling over N GPUs: tools/gpu_c4n.sh <tag> N
OUT=gpurun_out; mkdir -p $OUT; TAG=$1; N=$2
S0=$SECONDS
if [ "$N" = "1" ]; then
  timeout 1200 python bench.py --gpus 1 --config c4 --steps 2 --warmup 3 --no-cpu-baseline --strong > $OUT/${TAG}_n$N.json 2> $OUT/${TAG}_n$N.err
else
  timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2971$N bench.py --gpus $N --config c4 --steps 2 --warmup 3 --strong > $OUT/${TAG}_n$N.json 2> $OUT/${TAG}_n$N.err
fi
echo "N=$N rc=$? wall=$((SECONDS-S0))s"
python - <<PY
import json
try:
    d=json.loads([l for l in open("$OUT/${TAG}_n$N.json") if l.startswith("{")][-1])
    print("  value %.4g e2e %.4g ms/step %.1f frac %.3f path %s | %s" % (d["value"], d["e2e"]["value"], d["ms_per_step"], d["roofline"]["frac"], d["config"]["path"], d["config"]["parallelism"]))
except Exception as e:
    print("  failed", e); print(open("$OUT/${TAG}_n$N.err").read()[-1500:])
PY
