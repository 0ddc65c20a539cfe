#!/bin/bash
# On the GPU box: turn an .ncu-rep into small CSV pages (raw metrics + per-instruction source view) and drop the report
# (gpurun_out/ is capped at 64 MiB).  usage: tools/ncu_export.sh gpurun_out/name   (without .ncu-rep)
for b in "$@"; do
  [ -f $b.ncu-rep ] || { echo "missing $b.ncu-rep"; continue; }
  ncu -i $b.ncu-rep --page raw --csv > $b.raw.csv 2>/dev/null
  ncu -i $b.ncu-rep --page source --csv --print-source sass > $b.sass.csv 2>/dev/null
  ncu -i $b.ncu-rep --page details --csv > $b.details.csv 2>/dev/null
  gzip -f $b.sass.csv
  rm -f $b.ncu-rep
  ls -la $b.*
done
