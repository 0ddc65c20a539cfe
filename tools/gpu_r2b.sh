#!/bin/bash
# round 2, visit b: GPU test suite on the new host code, the new bench line at full C5 size on one GPU, full-size ncu capture
OUT=gpurun_out; mkdir -p $OUT; T=r2b
SECONDS=0
timeout 1200 python -m pytest tests -m gpu -x -q > $OUT/${T}_pytest.log 2>&1; echo "pytest rc=$? wall ${SECONDS}s"; tail -4 $OUT/${T}_pytest.log
timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' > $OUT/${T}_smoke.log 2>&1; echo "smoke rc=$? wall ${SECONDS}s"; tail -2 $OUT/${T}_smoke.log
timeout 1500 python bench.py --steps 6 --warmup 3 --only-configs c1,c2,c2_reference_schedule_b128 --ab-tunings 1,0 > $OUT/${T}_bench_c5_full.json 2> $OUT/${T}_bench_c5_full.err; echo "bench rc=$? wall ${SECONDS}s"
cut -c1-1500 $OUT/${T}_bench_c5_full.json; tail -5 $OUT/${T}_bench_c5_full.err
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:accumulate_memo_kernel -s 300 -c 1 -o $OUT/${T}_memo_c5full -f \
   python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --no-configs > $OUT/${T}_ncu_memo_c5full.log 2>&1; echo "ncu memo c5 full rc=$? wall ${SECONDS}s"
tools/ncu_export.sh $OUT/${T}_memo_c5full
du -sh $OUT
