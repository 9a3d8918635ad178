#!/bin/bash
# round 2, first session on the box: the GPU test suite, the C2 bench line, the per-band probe and a launch list of one band
OUT=gpurun_out
mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -x -q --durations=15 > $OUT/r02a_pytest.log 2>&1; echo "pytest rc=$?"; tail -25 $OUT/r02a_pytest.log
timeout 600 python bench.py > $OUT/r02a_bench_C2.json 2> $OUT/r02a_bench.err; echo "bench rc=$?"; cut -c1-400 $OUT/r02a_bench_C2.json
timeout 600 python tools/band_probe.py C2 8 6 1 > $OUT/r02a_band_probe_C2_n8.jsonl 2> $OUT/r02a_band_probe.err; echo "probe rc=$?"; cat $OUT/r02a_band_probe_C2_n8.jsonl | cut -c1-700
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/r02a_launches_band.csv \
    python tools/tile_probe.py 465 69 > $OUT/r02a_ncu_band.log 2>&1; echo "ncu band rc=$?"
