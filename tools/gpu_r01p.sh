set -x
OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
SECONDS=0
timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/r01p_pytest.log 2>&1; echo "pytest rc=$? wall ${SECONDS}s"
tail -8 $OUT/r01p_pytest.log
tools/gpu_ab.sh r01p_c4 c4 0.25 ssvi_tf_b200/libssvi_b200_old.so ssvi_tf_b200/libssvi_b200.so ssvi_tf_b200/libssvi_b200_t4.so
for W in 256; do
  timeout 600 python bench.py --config c4 --scale 0.25 --steps 3 --warmup 3 --no-cpu-baseline --no-e2e --sweep-chunk $W > $OUT/r01p_c4_w$W.json 2> $OUT/r01p_c4_w$W.err
  python -c "
import json; d=json.load(open('$OUT/r01p_c4_w$W.json')); print('sweep-chunk $W value %.4g ms/step %.2f frac %.3f' % (d['value'], d['ms_per_step'], d['roofline']['frac']))"
done
timeout 900 ncu --set full --clock-control none --import-source on -k regex:accumulate_diag_lin -s 2 -c 1 -o $OUT/r01p_ncu_diag_lin -f \
   python bench.py --config c4 --scale 0.05 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > $OUT/r01p_ncu.log 2>&1
echo "ncu rc=$?"; tail -3 $OUT/r01p_ncu.log
