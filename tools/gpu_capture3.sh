#!/bin/bash
# GPU-box session for profiles/ (round 1, ray-segment kernels): bench (both arms), ncu launch list, ncu --set full of the
# kernels of one steady-state frame, and of the sphere-list segment kernel. usage (repo root, on the box): bash tools/gpu_capture3.sh <tag>
TAG=${1:-r01o}
OUT=gpurun_out
mkdir -p $OUT
python bench.py > $OUT/${TAG}_bench_C2.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > $OUT/${TAG}_bench_reference_C2.json 2> $OUT/${TAG}_bench_reference.err; echo "ref rc=$?"
[ -n "$VR_CAPTURE_LIGHT" ] || { python bench.py --eval-variant 2 --no-cpu-baseline > $OUT/${TAG}_bench_C2_exact_kernel.json 2>> $OUT/${TAG}_bench.err; echo "bench v2 rc=$?"; }
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $OUT/${TAG}_launches_C2.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on \
    -k 'regex:k_eval_seg|k_eval_point3|k_march|k_lightgen<|k_resolve|k_flag_near|k_build_seg|k_list_flagged' -s 36 -c 13 -f \
    -o $OUT/${TAG}_full_frame python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
ncu -i $OUT/${TAG}_full_frame.ncu-rep --page raw --csv > $OUT/${TAG}_ncu_full_frame_kernels.csv 2>/dev/null
[ -n "$VR_CAPTURE_LIGHT" ] && { cut -c1-300 $OUT/${TAG}_bench_C2.json; exit 0; }   # VR_CAPTURE_LIGHT=1: C2 frame only
python tools/eval_variants.py sphere 1000 32 8 > $OUT/${TAG}_sphere_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:k_eval_seg -s 2 -c 1 -f -o $OUT/${TAG}_full_sphere \
    python tools/eval_variants.py sphere 1000 32 8 > $OUT/${TAG}_ncu_sphere.log 2>&1; echo "ncu sphere rc=$?"
ncu -i $OUT/${TAG}_full_sphere.ncu-rep --page raw --csv > $OUT/${TAG}_ncu_full_k_eval_seg_sphere.csv 2>/dev/null
cut -c1-300 $OUT/${TAG}_bench_C2.json
