#!/bin/bash
# round 2, final captures on one GPU for profiles/: GPU test suite, both bench arms, the band probe, launch lists (bench run,
# full frame, one 1/8 band), ncu --set full of one steady-state frame, every BASELINE.json configuration.
# usage (repo root, on the box): bash tools/gpu_final_r02.sh <tag>
TAG=${1:-r02}
OUT=gpurun_out
mkdir -p $OUT
timeout 900 python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/${TAG}_pytest_gpu.log
python bench.py > $OUT/${TAG}_bench_C2.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?"; cut -c1-200 $OUT/${TAG}_bench_C2.json
python bench.py --impl reference --steps 3 --warmup 1 > $OUT/${TAG}_bench_reference_C2.json 2> $OUT/${TAG}_bench_reference.err; echo "ref rc=$?"; cut -c1-200 $OUT/${TAG}_bench_reference_C2.json
python tools/band_probe.py C2 8 6 1 > $OUT/${TAG}_band_probe_C2_n8.jsonl 2> $OUT/${TAG}_band_probe.err; echo "probe rc=$?"; tail -1 $OUT/${TAG}_band_probe_C2_n8.jsonl | cut -c1-300
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/${TAG}_launches_C2.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
for spec in "0 1080 full_frame" "467 68 band_467_68"; do
  set -- $spec
  ncu --metrics gpu__time_duration.sum,sm__cycles_active.avg,sm__cycles_elapsed.max --clock-control none -c 400 --csv --log-file $OUT/${TAG}_launches_C2_$3.csv \
    python tools/tile_probe.py $1 $2 > $OUT/${TAG}_ncu_$3.log 2>&1; echo "ncu $3 rc=$?"
done
ncu --set full --clock-control none --import-source on \
    -k 'regex:k_eval_seg|k_eval_point3|k_march|k_lightgen<|k_resolve|k_flag_near|k_build_seg|k_list_flagged' -s 33 -c 11 -f \
    -o $OUT/${TAG}_full_frame python tools/tile_probe.py 0 1080 > $OUT/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
ncu -i $OUT/${TAG}_full_frame.ncu-rep --page raw --csv > $OUT/${TAG}_ncu_full_frame_kernels.csv 2>/dev/null
python tools/ncu_summary.py $OUT/${TAG}_ncu_full_frame_kernels.csv > $OUT/${TAG}_ncu_summary.txt 2>&1; cut -c1-300 $OUT/${TAG}_ncu_summary.txt | head -20
rm -f $OUT/${TAG}_full_frame.ncu-rep
python tools/run_configs.py > $OUT/${TAG}_configs.jsonl 2> $OUT/${TAG}_configs.err; echo "configs rc=$?"
python - <<P
import json
for l in open("$OUT/${TAG}_configs.jsonl"):
    d=json.loads(l); print(d.get("config"), d.get("ms_per_frame"), d.get("parity"))
P
