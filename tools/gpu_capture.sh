#!/bin/bash
# One GPU-box session: parity tests, bench (both arms), ncu launch list, ncu --set full of the hot kernel.
# usage (from the repo root, on the box): bash tools/gpu_capture.sh <tag> [kernel-regex]
TAG=${1:-r01}
KREGEX=${2:-k_eval_point2}
OUT=gpurun_out
mkdir -p $OUT
python -m pytest tests -m gpu -x -q > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?" | tee -a $OUT/${TAG}_pytest.log
python bench.py > $OUT/${TAG}_bench.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?"
python bench.py --impl reference --steps 3 --warmup 1 > $OUT/${TAG}_bench_reference.json 2> $OUT/${TAG}_bench_reference.err; echo "ref rc=$?"
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $OUT/${TAG}_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_ncu_launches.log 2>&1; echo "ncu launches rc=$?"
ncu --set full --clock-control none --import-source on -k regex:${KREGEX} -s 3 -c 1 -f -o $OUT/${TAG}_full_${KREGEX} \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $OUT/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
ncu -i $OUT/${TAG}_full_${KREGEX}.ncu-rep --page raw --csv > $OUT/${TAG}_full_${KREGEX}_raw.csv 2>/dev/null
tail -3 $OUT/${TAG}_pytest.log; cat $OUT/${TAG}_bench.json
