#!/bin/bash
# round 2 session: GPU test suite, C2 bench line, per-band probe (what one of 8 GPUs spends on its band)
TAG=${1:-r02b}
OUT=gpurun_out
mkdir -p $OUT
timeout 1500 python -m pytest tests -m gpu -x -q --durations=10 > $OUT/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -18 $OUT/${TAG}_pytest.log
timeout 600 python bench.py > $OUT/${TAG}_bench_C2.json 2> $OUT/${TAG}_bench.err; echo "bench rc=$?"; cut -c1-300 $OUT/${TAG}_bench_C2.json; python - <<P
import json
d=json.loads(open("$OUT/${TAG}_bench_C2.json").read().strip().splitlines()[-1])
print({k:d[k] for k in ("ms_per_step","passes_ms_per_step","gpu_launches")}, d["e2e"]["ms_per_step"], d["roofline"]["frac"])
P
timeout 600 python tools/band_probe.py C2 8 6 1 > $OUT/${TAG}_band_probe_C2_n8.jsonl 2> $OUT/${TAG}_band_probe.err; echo "probe rc=$?"; cut -c1-420 $OUT/${TAG}_band_probe_C2_n8.jsonl
