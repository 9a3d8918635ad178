#!/bin/bash
# GPU-box session for profiles/: bench (both arms), ncu launch list, ncu --set full of every kernel of one frame,
# and of the packed sphere kernel. usage (repo root, on the box): bash tools/gpu_capture2.sh <tag>
TAG=${1:-r01h}
OUT=gpurun_out
mkdir -p $OUT
python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > $OUT/${TAG}_bench_reference.json 2> $OUT/${TAG}_bench_reference.err; echo "ref rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k 'regex:k_eval_point3|k_march|k_lightgen<|k_resolve' -s 18 -c 6 -f \
    -o $OUT/${TAG}_full_frame python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
ncu -i $OUT/${TAG}_full_frame.ncu-rep --page raw --csv > $OUT/${TAG}_full_frame_raw.csv 2>/dev/null
python tools/eval_variants.py sphere 1000 32 2 > $OUT/${TAG}_sphere_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_eval_sphere2 -s 2 -c 1 -f -o $OUT/${TAG}_full_sphere \
    python tools/eval_variants.py sphere 1000 32 2 > $OUT/${TAG}_ncu_sphere.log 2>&1; echo "ncu sphere rc=$?"
ncu -i $OUT/${TAG}_full_sphere.ncu-rep --page raw --csv > $OUT/${TAG}_full_sphere_raw.csv 2>/dev/null
cat $OUT/${TAG}_bench.json | cut -c1-400
