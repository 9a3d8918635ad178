#!/bin/bash
# round 2, one 8-GPU box, short re-run after the host-side changes of the end-to-end path: C2 at N = 1, 2, 4, 8, C3 at
# N = 8, the C++ host on C2 at 8 GPUs. usage (repo root): bash tools/gpu_scaling3.sh <tag>
TAG=${1:-r02g}
OUT=gpurun_out
mkdir -p $OUT
run() {  # workload N extra...
  WL=$1; N=$2; shift 2
  if [ $N = 1 ]; then
    python bench.py --gpus 1 --workload $WL --no-cpu-baseline "$@" > $OUT/${TAG}_bench_${WL}_n1.json 2> $OUT/${TAG}_bench_${WL}_n1.err
  else
    python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) \
        bench.py --gpus $N --workload $WL "$@" > $OUT/${TAG}_bench_${WL}_n$N.json 2> $OUT/${TAG}_bench_${WL}_n$N.err
  fi
  echo "$WL n$N rc=$?"
  tail -1 $OUT/${TAG}_bench_${WL}_n$N.json | python -c "
import sys,json
d=json.loads(sys.stdin.read()); e=d['e2e']; print('$WL', d['n_gpus'], round(d['ms_per_step'],4), '%.4g'%d['value'], 'e2e', round(e['ms_per_step'],4), e.get('passes_ms_per_step_rank0'), d.get('multi_gpu_image_equals_single_gpu'))"
}
for N in 8 1 2 4; do run C2 $N; done
run C3 8 --steps 8 --warmup 3
H=volumerenderer_b200/vr_headless
$H --grid cloud:256 --algo point --size 1920x1080 --set maxLights=1000 --lightgen 32 --gpus 8 --frames 4 --bench 20 > $OUT/${TAG}_headless_C2_n8.json 2> $OUT/${TAG}_headless_C2_n8.err; echo "headless C2 n8 rc=$?"; cut -c1-400 $OUT/${TAG}_headless_C2_n8.json
