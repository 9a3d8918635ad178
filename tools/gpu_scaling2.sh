#!/bin/bash
# round 2, one 8-GPU box: strong scaling of C2 and C3 (bench.py under torchrun at N = 1, 2, 4, 8), the C++ multi-GPU host
# (vr_headless --gpus 8) on C2 next to it, and the sample-parallel progressive run of C4-A (64 frames) at 1 vs 8 GPUs.
# usage (repo root): bash tools/gpu_scaling2.sh <tag>
TAG=${1:-r02}
OUT=gpurun_out
mkdir -p $OUT
for WL in C2 C3; do
  EXTRA=""; [ $WL = C3 ] && EXTRA="--steps 8 --warmup 3 --cpu-seconds 6"
  python bench.py --gpus 1 --workload $WL $EXTRA > $OUT/${TAG}_bench_${WL}_n1.json 2> $OUT/${TAG}_bench_${WL}_n1.err; echo "$WL n1 rc=$?"
  for N in 2 4 8; do
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) \
        bench.py --gpus $N --workload $WL $EXTRA > $OUT/${TAG}_bench_${WL}_n$N.json 2> $OUT/${TAG}_bench_${WL}_n$N.err; echo "$WL n$N rc=$?"
  done
  for N in 1 2 4 8; do tail -1 $OUT/${TAG}_bench_${WL}_n$N.json | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print('$WL', d['n_gpus'], round(d['ms_per_step'],4), '%.4g'%d['value'], 'e2e', round(d['e2e']['ms_per_step'],4), d.get('multi_gpu_image_equals_single_gpu'))"; done
done
H=volumerenderer_b200/vr_headless
C2ARGS="--grid cloud:256 --algo point --size 1920x1080 --set maxLights=1000 --lightgen 32"
for N in 1 8; do
  G="--gpus $N"; [ $N = 1 ] && G="--devices 0"
  $H $C2ARGS $G --frames 4 --bench 20 > $OUT/${TAG}_headless_C2_n$N.json 2> $OUT/${TAG}_headless_C2_n$N.err; echo "headless C2 n$N rc=$?"; cut -c1-400 $OUT/${TAG}_headless_C2_n$N.json
done
C4ARGS="--grid cloud:512 --algo beam --size 1920x1080 --set maxLights=256 --lightgen 8 --frames 64"
for N in 1 8; do
  G="--gpus $N"; [ $N = 1 ] && G="--devices 0"
  $H $C4ARGS $G --progressive 1 --pfm $OUT/${TAG}_c4a_n$N.pfm > $OUT/${TAG}_headless_C4A_progressive_n$N.json 2> $OUT/${TAG}_headless_C4A_n$N.err; echo "headless C4-A progressive n$N rc=$?"; cut -c1-400 $OUT/${TAG}_headless_C4A_progressive_n$N.json
done
$H $C4ARGS --gpus 8 > $OUT/${TAG}_headless_C4A_tiled_n8.json 2>> $OUT/${TAG}_headless_C4A_n8.err; echo "headless C4-A tiled n8 rc=$?"; cut -c1-400 $OUT/${TAG}_headless_C4A_tiled_n8.json
python - <<P
import numpy as np
def pfm(p):
    f=open(p,'rb'); t=f.readline(); w,h=map(int,f.readline().split()); s=float(f.readline()); a=np.frombuffer(f.read(),'<f4' if s<0 else '>f4'); return a.reshape(h,w,-1)
a,b=pfm("$OUT/${TAG}_c4a_n1.pfm"),pfm("$OUT/${TAG}_c4a_n8.pfm")
d=np.abs(a-b); print("C4-A 64 frames, sample-parallel 8 GPUs vs 1 GPU: max abs diff %.3g, max rel %.3g, mean %.4g" % (d.max(), (d/np.maximum(np.abs(a),1e-6)).max(), a.mean()))
P
rm -f $OUT/${TAG}_c4a_n*.pfm
