#!/bin/bash
# round 2, visit a: C5 profile of the shipped build (launch list, full ncu captures with source, exported to CSV on the box)
OUT=gpurun_out; mkdir -p $OUT; T=r2a
SECONDS=0
A="--config c5 --scale 0.2 --steps 1 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 1200 -c 500 --csv --log-file $OUT/${T}_launches_c5s02.csv \
   python bench.py $A > $OUT/${T}_ncu_list.log 2>&1; echo "ncu list rc=$? wall ${SECONDS}s"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:accumulate_memo_kernel -s 500 -c 2 -o $OUT/${T}_memo_c5 -f \
   python bench.py $A > $OUT/${T}_ncu_memo_c5.log 2>&1; echo "ncu memo c5 rc=$? wall ${SECONDS}s"
tools/ncu_export.sh $OUT/${T}_memo_c5
timeout 900 ncu --set full --clock-control none --import-source on -k regex:row_update_full_kernel -s 4 -c 1 -o $OUT/${T}_rowupd_c5 -f \
   python bench.py $A > $OUT/${T}_ncu_rowupd_c5.log 2>&1; echo "ncu rowupd c5 rc=$? wall ${SECONDS}s"
tools/ncu_export.sh $OUT/${T}_rowupd_c5
timeout 600 ncu --set full --clock-control none --import-source on -k regex:accumulate_memo_kernel -s 40 -c 1 -o $OUT/${T}_memo_c2 -f \
   python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e > $OUT/${T}_ncu_memo_c2.log 2>&1; echo "ncu memo c2 rc=$? wall ${SECONDS}s"
tools/ncu_export.sh $OUT/${T}_memo_c2
du -sh $OUT
