set -x
OUT=gpurun_out; mkdir -p $OUT
tools/test_exp | tee $OUT/r01o_test_exp.txt
SECONDS=0
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/r01o_pytest.log 2>&1; echo "pytest rc=$? wall ${SECONDS}s"
tail -4 $OUT/r01o_pytest.log
tools/gpu_ab.sh r01o_ab c2 1 ssvi_tf_b200/libssvi_b200_old.so ssvi_tf_b200/libssvi_b200.so
