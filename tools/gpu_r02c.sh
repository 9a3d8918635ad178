#!/bin/bash
# round 2: launch list of ONE 1/8-frame band of C2 (kernel durations + SM-active share = how much of each launch is tail)
OUT=gpurun_out
mkdir -p $OUT
python tools/tile_probe.py 467 68 > $OUT/r02c_tile.log 2>&1; echo "tile rc=$?"; tail -3 $OUT/r02c_tile.log
timeout 900 ncu --metrics gpu__time_duration.sum,sm__cycles_active.avg,sm__cycles_elapsed.max,launch__grid_size --clock-control none -c 600 --csv --log-file $OUT/r02c_launches_band.csv \
    python tools/tile_probe.py 467 68 > $OUT/r02c_ncu_band.log 2>&1; echo "ncu band rc=$?"
