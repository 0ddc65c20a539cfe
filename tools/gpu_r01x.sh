set -x
OUT=gpurun_out; mkdir -p $OUT
SECONDS=0
timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/r01x_pytest.log 2>&1; echo "pytest rc=$? wall ${SECONDS}s"; tail -4 $OUT/r01x_pytest.log
tools/gpu_ab.sh r01x_c4 c4 0.25 ssvi_tf_b200/libssvi_b200.so
tools/gpu_ab.sh r01x_c4d c4d 0.25 ssvi_tf_b200/libssvi_b200.so
