#!/bin/bash
# quick N-GPU check of the C2 bench line (device-resident and end to end): bash tools/gpu_n2_quick.sh <tag> <N>
TAG=${1:-r02q}; N=${2:-2}
OUT=gpurun_out
mkdir -p $OUT
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29502 \
    bench.py --gpus $N --workload C2 > $OUT/${TAG}_bench_C2_n$N.json 2> $OUT/${TAG}_bench_C2_n$N.err; echo "C2 n$N rc=$?"
tail -1 $OUT/${TAG}_bench_C2_n$N.json | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print({k:d.get(k) for k in ('ms_per_step','multi_gpu_image_equals_single_gpu')}); print(d['per_rank_ms_per_step']); print(d['e2e']); print(d['config']['bands'])"
