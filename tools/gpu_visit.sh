#!/bin/bash
# One GPU-box visit, parameterised (replaces the per-visit one-off scripts).  Runs under gpurun from the repo root:
#   gpurun --timeout 1800 -- 'bash tools/gpu_visit.sh <tag> <step> [<step> ...]'
# Steps (each writes gpurun_out/<tag>_*; quote a step that carries arguments):
#   suite                      python -m pytest tests -m gpu
#   multi                      python -m pytest tests/test_gpu_multi.py (needs >= 2 GPUs)
#   smoke                      __graft_entry__.smoke()
#   "bench <name> <args>"      python bench.py <args>                      -> <tag>_bench_<name>.json
#   "dist <n> <name> <args>"   torchrun --nproc-per-node n bench.py --gpus n <args>
#   "launches <name> <skip> <count> <cmd...>"   ncu launch list of <cmd> (run plain first)  -> <tag>_launches_<name>.csv + summary
#   "klaunches <name> <kernel regex> <skip> <count> <cmd...>"   the same, counting only kernels whose name matches
#   "ncu <name> <kernel regex> <skip> <count> <cmd...>"   ncu --set full with source of the matching kernel, exported to
#                              CSV pages on the box (tools/ncu_export.sh; gpurun_out/ is capped at 64 MiB)
#   "factorize [env...]"       tools/factorize_timing.py
TAG=$1; shift
OUT=gpurun_out; mkdir -p $OUT; SECONDS=0
for step in "$@"; do
  set -- $step; kind=$1; shift
  case $kind in
    suite) timeout 1800 python -m pytest tests -m gpu -q > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "suite rc=$?"; tail -4 $OUT/${TAG}_pytest_gpu.log;;
    multi) timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -v > $OUT/${TAG}_pytest_multi.log 2>&1; echo "multi rc=$?"; tail -6 $OUT/${TAG}_pytest_multi.log;;
    smoke) timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' > $OUT/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $OUT/${TAG}_smoke.log;;
    bench) name=$1; shift; timeout 1500 python bench.py "$@" > $OUT/${TAG}_bench_$name.json 2> $OUT/${TAG}_bench_$name.err; echo "bench $name rc=$?"
           python tools/bench_line.py $OUT/${TAG}_bench_$name.json; tail -3 $OUT/${TAG}_bench_$name.err;;
    dist) n=$1; name=$2; shift 2
          timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 \
            bench.py --gpus $n "$@" > $OUT/${TAG}_bench_$name.json 2> $OUT/${TAG}_bench_$name.err; echo "bench $name (N=$n) rc=$?"
          python tools/bench_line.py $OUT/${TAG}_bench_$name.json; tail -3 $OUT/${TAG}_bench_$name.err;;
    launches) name=$1; skip=$2; count=$3; shift 3
          timeout 900 "$@" > $OUT/${TAG}_plain_$name.log 2>&1 &&
          timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -s $skip -c $count --csv \
            --log-file $OUT/${TAG}_launches_$name.csv "$@" > $OUT/${TAG}_ncu_$name.log 2>&1; echo "launches $name rc=$?"
          python tools/launch_summary.py $OUT/${TAG}_launches_$name.csv | tee $OUT/${TAG}_launches_${name}_summary.txt | head -14;;
    klaunches) name=$1; regex=$2; skip=$3; count=$4; shift 4       # launch list restricted to kernels matching <regex>
          timeout 900 "$@" > $OUT/${TAG}_plain_$name.log 2>&1 &&
          timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:$regex -s $skip -c $count --csv \
            --log-file $OUT/${TAG}_launches_$name.csv "$@" > $OUT/${TAG}_ncu_$name.log 2>&1; echo "klaunches $name rc=$?"
          python tools/launch_summary.py $OUT/${TAG}_launches_$name.csv | tee $OUT/${TAG}_launches_${name}_summary.txt | head -14;;
    ncu) name=$1; regex=$2; skip=$3; count=$4; shift 4
          timeout 900 "$@" > $OUT/${TAG}_plain_$name.log 2>&1 &&
          timeout 1500 ncu --set full --clock-control none --import-source on -k regex:$regex -s $skip -c $count \
            -o $OUT/${TAG}_$name -f "$@" > $OUT/${TAG}_ncu_$name.log 2>&1; echo "ncu $name rc=$?"
          tools/ncu_export.sh $OUT/${TAG}_$name;;
    factorize) env "$@" timeout 300 python tools/factorize_timing.py > $OUT/${TAG}_factorize$(echo "$@" | tr -d ' =').json 2>> $OUT/${TAG}_factorize.err
          echo "factorize rc=$?"; cat $OUT/${TAG}_factorize$(echo "$@" | tr -d ' =').json | cut -c1-700;;
    *) echo "unknown step: $kind";;
  esac
  echo "-- $kind done at ${SECONDS}s"
done
du -sh $OUT
