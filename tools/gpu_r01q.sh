set -x
OUT=gpurun_out; mkdir -p $OUT
for lib in libssvi_b200 libssvi_b200_e2t8pf0 libssvi_b200_e4t4pf0 libssvi_b200_e4t2pf0; do
  SSVI_B200_LIB=$PWD/ssvi_tf_b200/$lib.so timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "sweep_diag" > $OUT/r01q_pytest_$lib.log 2>&1; echo "$lib pytest rc=$?"; tail -2 $OUT/r01q_pytest_$lib.log
done
tools/gpu_ab.sh r01q_c4 c4 0.25 ssvi_tf_b200/libssvi_b200.so ssvi_tf_b200/libssvi_b200_e2t8pf0.so ssvi_tf_b200/libssvi_b200_e4t4pf0.so ssvi_tf_b200/libssvi_b200_e4t2pf0.so
