set -x
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python -m pytest tests/test_gpu_parity.py tests/test_gpu_api.py -m gpu -x -q -k "subsampled or minibatch or drop_in or factorize" > $OUT/r01y_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 $OUT/r01y_pytest.log
timeout 600 python bench.py --batch-size 128 --steps 5 --warmup 3 --no-cpu-baseline > $OUT/r01y_c2_sub128.json 2> $OUT/r01y_c2_sub128.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('$OUT/r01y_c2_sub128.json')); print('subsampled B=128 value %.4g e2e %.4g ms/step %.2f kernel_ms %.3f share %.3f launches %d' % (d['value'], d['e2e']['value'], d['ms_per_step'], d['roofline']['kernel_ms_per_launch'], d['roofline']['kernel_share_of_step'], d['gpu_launches']))"
tail -3 $OUT/r01y_c2_sub128.err
