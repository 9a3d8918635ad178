#!/bin/bash
# round 2: BASELINE.json's C5 (3840x2160 through the 1024^3 cloud, sphere lights + path mode, "image-tiled across 8 B200") as
# eight balanced row bands rendered one after the other on ONE GPU (what each of eight GPUs would do), with the bands of one
# frame compared bit for bit against the whole-frame render; then the round-end checks of the final build (smoke, both bench arms).
# usage (repo root, on the box): bash tools/gpu_c5_tiled.sh <tag>
TAG=${1:-r02h}
OUT=gpurun_out
mkdir -p $OUT
for WL in C5S C5P; do
  timeout 300 python tools/band_probe.py $WL 8 4 1 1 > $OUT/${TAG}_band_probe_${WL}_n8.jsonl 2> $OUT/${TAG}_band_probe_${WL}.err; echo "probe $WL rc=$?"
  head -1 $OUT/${TAG}_band_probe_${WL}_n8.jsonl | cut -c1-300; tail -2 $OUT/${TAG}_band_probe_${WL}_n8.jsonl | cut -c1-700
done
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > $OUT/${TAG}_smoke.log 2>&1; echo "smoke rc=$?"; tail -2 $OUT/${TAG}_smoke.log
python bench.py > $OUT/${TAG}_bench_C2.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?"; cut -c1-260 $OUT/${TAG}_bench_C2.json
python bench.py --impl reference --steps 2 --warmup 1 > $OUT/${TAG}_bench_reference_C2.json 2> $OUT/${TAG}_bench_reference.err; echo "ref rc=$?"; cut -c1-200 $OUT/${TAG}_bench_reference_C2.json
