set -x
OUT=gpurun_out; mkdir -p $OUT
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv
SECONDS=0
timeout 1500 python -m pytest tests -m gpu -x -q > $OUT/r01n_pytest.log 2>&1; echo "pytest rc=$? wall ${SECONDS}s"
tail -8 $OUT/r01n_pytest.log
timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' > $OUT/r01n_smoke.log 2>&1; echo "smoke rc=$?"; tail -3 $OUT/r01n_smoke.log
tools/gpu_ab.sh r01n_ab c2 1 ssvi_tf_b200/libssvi_b200.so ssvi_tf_b200/libssvi_b200_pf2_b4.so ssvi_tf_b200/libssvi_b200_pf1_b3.so ssvi_tf_b200/libssvi_b200_pf2_b3.so ssvi_tf_b200/libssvi_b200_pf3_b3.so
tools/gpu_ab.sh r01n_c4 c4 0.25 ssvi_tf_b200/libssvi_b200.so
