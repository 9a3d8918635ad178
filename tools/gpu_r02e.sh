#!/bin/bash
# launch list (duration + SM-active share) of the full C2 frame and of one 1/8 band
OUT=gpurun_out
TAG=${1:-r02e}
mkdir -p $OUT
for spec in "0 1080 full" "467 68 band"; do
  set -- $spec
  timeout 900 ncu --metrics gpu__time_duration.sum,sm__cycles_active.avg,sm__cycles_elapsed.max --clock-control none -c 400 --csv --log-file $OUT/${TAG}_launches_$3.csv \
    python tools/tile_probe.py $1 $2 > $OUT/${TAG}_ncu_$3.log 2>&1; echo "ncu $3 rc=$?"
done
