#!/bin/bash
# round 2, one 8-GPU box: BASELINE.json's C5 (3840x2160, 1024^3 cloud) on eight GPUs from the C++ host: sphere lights as
# image bands (device time per frame), path mode as the sample-parallel progressive run (wall time of 16 frames).
# usage (repo root, on the box): bash tools/gpu_c5_n8.sh <tag>
TAG=${1:-r02i}
OUT=gpurun_out
mkdir -p $OUT
H=volumerenderer_b200/vr_headless
C5="--grid cloud:1024 --size 3840x2160 --set maxLights=1000 --lightgen 32"
timeout 45 $H $C5 --algo sphere --gpus 8 --balance-rounds 4 --frames 2 --bench 6 > $OUT/${TAG}_headless_C5S_n8.json 2> $OUT/${TAG}_headless_C5S_n8.err; echo "C5 sphere bands n8 rc=$?"; cut -c1-500 $OUT/${TAG}_headless_C5S_n8.json
timeout 35 $H $C5 --algo path --gpus 8 --frames 16 --progressive 1 > $OUT/${TAG}_headless_C5P_progressive_n8.json 2> $OUT/${TAG}_headless_C5P_n8.err; echo "C5 path sample-parallel n8 rc=$?"; cut -c1-400 $OUT/${TAG}_headless_C5P_progressive_n8.json
