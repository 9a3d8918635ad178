#!/bin/bash
# round 2: ncu --set full of the kernels of one steady-state C2 frame (and optionally one 1/8 band), summaries for profiles/
# usage (repo root, on the box): bash tools/gpu_capture4.sh <tag>
TAG=${1:-r02}
OUT=gpurun_out
mkdir -p $OUT
ncu --set full --clock-control none --import-source on \
    -k 'regex:k_eval_seg|k_eval_point3|k_march|k_lightgen<|k_resolve|k_flag_near|k_build_seg|k_list_flagged' -s 33 -c 11 -f \
    -o $OUT/${TAG}_full_frame python tools/tile_probe.py 0 1080 > $OUT/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
ncu -i $OUT/${TAG}_full_frame.ncu-rep --page raw --csv > $OUT/${TAG}_ncu_full_frame_kernels.csv 2>/dev/null
python tools/ncu_summary.py $OUT/${TAG}_ncu_full_frame_kernels.csv | cut -c1-420
