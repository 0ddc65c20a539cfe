set -x
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "sweep_diag" > $OUT/r01t_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/r01t_pytest.log
tools/gpu_ab.sh r01t_c4d c4d 0.25 ssvi_tf_b200/libssvi_b200_old.so ssvi_tf_b200/libssvi_b200.so
timeout 600 python tools/factorize_timing.py > $OUT/r01t_factorize_timing.json 2> $OUT/r01t_factorize_timing.err; echo "timing rc=$?"; cat $OUT/r01t_factorize_timing.json; tail -3 $OUT/r01t_factorize_timing.err
