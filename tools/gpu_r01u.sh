set -x
OUT=gpurun_out; mkdir -p $OUT
SECONDS=0
timeout 900 python bench.py --config c4 --steps 3 --warmup 3 --no-cpu-baseline > $OUT/r01u_c4_full.json 2> $OUT/r01u_c4_full.err; echo "full c4 rc=$? wall ${SECONDS}s"
python -c "
import json; d=json.load(open('$OUT/r01u_c4_full.json')); print('C4 full value %.4g e2e %.4g ms/step %.2f frac %.3f launches %d' % (d['value'], d['e2e']['value'], d['ms_per_step'], d['roofline']['frac'], d['gpu_launches']))"
SECONDS=0
timeout 1200 python bench.py --config c3 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > $OUT/r01u_c3_full.json 2> $OUT/r01u_c3_full.err; echo "full c3 rc=$? wall ${SECONDS}s"
python -c "
import json; d=json.load(open('$OUT/r01u_c3_full.json')); print('C3 full value %.4g ms/step %.2f frac %.3f' % (d['value'], d['ms_per_step'], d['roofline']['frac']))"
tail -3 $OUT/r01u_c3_full.err
