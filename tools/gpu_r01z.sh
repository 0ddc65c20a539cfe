set -x
OUT=gpurun_out; mkdir -p $OUT
timeout 600 python bench.py --config c3 --scale 0.05 --steps 2 --warmup 2 --no-cpu-baseline --no-e2e > $OUT/r01z_c3_s05.json 2> $OUT/r01z_c3_s05.err; echo "bench rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 200 -c 400 --csv --log-file $OUT/r01z_launches_c3.csv \
  python bench.py --config c3 --scale 0.05 --steps 1 --warmup 1 --no-cpu-baseline --no-e2e > $OUT/r01z_ncu_list.log 2>&1; echo "ncu rc=$?"
python tools/launch_summary.py $OUT/r01z_launches_c3.csv | head -14
timeout 600 python bench.py --config c1 --steps 20 --warmup 5 --no-cpu-baseline > $OUT/r01z_c1.json 2> $OUT/r01z_c1.err; echo "c1 rc=$?"
python -c "
import json
for f in ['r01z_c3_s05','r01z_c1']:
    d=json.load(open('$OUT/%s.json'%f)); print(f,'value %.4g ms/step %.3f frac %.3f launches %d' % (d['value'], d['ms_per_step'], d['roofline']['frac'], d['gpu_launches']))"
