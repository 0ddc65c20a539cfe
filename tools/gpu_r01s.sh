set -x
OUT=gpurun_out; mkdir -p $OUT
SSVI_B200_LIB=$PWD/ssvi_tf_b200/libssvi_b200_new.so timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "sweep_diag" > $OUT/r01s_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $OUT/r01s_pytest.log
tools/gpu_ab.sh r01s_c4 c4 0.25 ssvi_tf_b200/libssvi_b200.so ssvi_tf_b200/libssvi_b200_new.so
SECONDS=0
SSVI_B200_LIB=$PWD/ssvi_tf_b200/libssvi_b200_new.so timeout 900 python bench.py --config c4 --steps 3 --warmup 3 --no-cpu-baseline --sweep-chunk 256 > $OUT/r01s_c4_full.json 2> $OUT/r01s_c4_full.err; echo "full c4 rc=$? wall ${SECONDS}s"
python -c "
import json; d=json.load(open('$OUT/r01s_c4_full.json')); print('C4 full value %.4g e2e %.4g ms/step %.2f frac %.3f' % (d['value'], d['e2e']['value'], d['ms_per_step'], d['roofline']['frac']))"
