set -x
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "sweep_diag" > $OUT/r01r_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 $OUT/r01r_pytest.log
tools/gpu_ab.sh r01r_c4 c4 0.25 ssvi_tf_b200/libssvi_b200_e4t4pf0.so ssvi_tf_b200/libssvi_b200.so
timeout 600 python bench.py --config c4 --scale 0.25 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --sweep-chunk 256 > $OUT/r01r_c4_w256.json 2> $OUT/r01r_c4_w256.err
python -c "
import json; d=json.load(open('$OUT/r01r_c4_w256.json')); print('sweep-chunk 256 value %.4g ms/step %.2f frac %.3f' % (d['value'], d['ms_per_step'], d['roofline']['frac']))"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:accumulate_diag_lin -s 2 -c 1 -o $OUT/r01r_ncu_diag_lin -f \
   python bench.py --config c4 --scale 0.05 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > $OUT/r01r_ncu.log 2>&1
echo "ncu rc=$?"
