set -x
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_fullsize.py -m gpu -x -q -k "sweep_table or tiled or full_size or config3" > $OUT/r01w_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/r01w_pytest.log
tools/gpu_ab.sh r01w_c2 c2 1 ssvi_tf_b200/libssvi_b200_prev.so ssvi_tf_b200/libssvi_b200.so ssvi_tf_b200/libssvi_b200_prev.so ssvi_tf_b200/libssvi_b200.so
