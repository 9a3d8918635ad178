#!/bin/bash
# two-GPU check of the N > 1 paths: bench C2 and C3 under torchrun, the C++ group (vr_headless --gpus 2, pytest group tests)
TAG=${1:-r02n2}
OUT=gpurun_out
mkdir -p $OUT
for WL in C2 C3; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29502 \
    bench.py --gpus 2 --workload $WL --steps 10 --warmup 3 > $OUT/${TAG}_bench_${WL}_n2.json 2> $OUT/${TAG}_bench_${WL}_n2.err; echo "$WL n2 rc=$?"
tail -1 $OUT/${TAG}_bench_${WL}_n2.json | python -c "
import sys,json
d=json.loads(sys.stdin.read()); print({k:d.get(k) for k in ('ms_per_step','value','multi_gpu_image_equals_single_gpu','per_rank_ms_per_step')}); print(d['e2e'])"
done
timeout 900 python -m pytest tests/test_gpu_group.py tests/test_gpu_headless.py -m gpu -x -q > $OUT/${TAG}_pytest_group.log 2>&1; echo "pytest rc=$?"; tail -3 $OUT/${TAG}_pytest_group.log
