set -x
OUT=gpurun_out; mkdir -p $OUT
SECONDS=0
timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/r02c_pytest.log 2>&1; echo "pytest rc=$? wall ${SECONDS}s"; tail -4 $OUT/r02c_pytest.log
timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' > $OUT/r02c_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $OUT/r02c_smoke.log
SECONDS=0
timeout 1200 python bench.py --config c3 --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > $OUT/r02c_c3_full.json 2> $OUT/r02c_c3_full.err; echo "full c3 rc=$? wall ${SECONDS}s"
python -c "
import json; d=json.load(open('$OUT/r02c_c3_full.json')); print('C3 full value %.4g ms/step %.2f frac %.3f' % (d['value'], d['ms_per_step'], d['roofline']['frac']))"
