set -x
OUT=gpurun_out; mkdir -p $OUT
SECONDS=0
timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/r02e_pytest.log 2>&1; echo "pytest rc=$? wall ${SECONDS}s"; tail -4 $OUT/r02e_pytest.log
timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' > $OUT/r02e_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $OUT/r02e_smoke.log
tools/gpu_ab.sh r02e_c2 c2 1 ssvi_tf_b200/libssvi_b200.so
tools/gpu_ab.sh r02e_c3 c3 0.05 ssvi_tf_b200/libssvi_b200.so
