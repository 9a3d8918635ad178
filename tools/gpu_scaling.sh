#!/bin/bash
# strong-scaling bench lines for profiles/: usage (repo root, on a box with 8 GPUs): bash tools/gpu_scaling.sh <tag>
TAG=${1:-r01m}
OUT=gpurun_out
mkdir -p $OUT
python bench.py --gpus 1 --no-cpu-baseline > $OUT/${TAG}_bench_C2_n1.json 2> $OUT/${TAG}_scaling.err; echo "n1 rc=$?"
for N in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500 + N)) \
      bench.py --gpus $N > $OUT/${TAG}_bench_C2_n$N.json 2>> $OUT/${TAG}_scaling.err; echo "n$N rc=$?"
done
for N in 1 2 4 8; do tail -1 $OUT/${TAG}_bench_C2_n$N.json | cut -c1-220; done
