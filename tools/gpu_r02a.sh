set -x
OUT=gpurun_out; mkdir -p $OUT
SECONDS=0
timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/r02a_pytest.log 2>&1; echo "pytest rc=$? wall ${SECONDS}s"; tail -6 $OUT/r02a_pytest.log
tools/gpu_ab.sh r02a_c3 c3 0.05 ssvi_tf_b200/libssvi_b200_prev.so ssvi_tf_b200/libssvi_b200.so
