set -x
OUT=gpurun_out; mkdir -p $OUT
SECONDS=0
SSVI_B200_LIB=$PWD/ssvi_tf_b200/libssvi_b200_new.so timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/r02d_pytest.log 2>&1; echo "pytest rc=$? wall ${SECONDS}s"; tail -4 $OUT/r02d_pytest.log
tools/gpu_ab.sh r02d_c2 c2 1 ssvi_tf_b200/libssvi_b200.so ssvi_tf_b200/libssvi_b200_new.so ssvi_tf_b200/libssvi_b200.so ssvi_tf_b200/libssvi_b200_new.so
