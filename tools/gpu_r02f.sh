#!/bin/bash
# segment-kernel configurations on the full C2 frame: eval_ms per variant (8: 256x2, 9: 128x3 pipelined, 10: 256x2 pipelined, 11: 128x3)
OUT=gpurun_out
mkdir -p $OUT
for v in 8 9 10 11; do
  echo "variant $v"; python tools/tile_probe.py 0 1080 $v 2>&1 | tail -2
done | tee $OUT/r02f_variants.log
for v in 8 9; do
  echo "band variant $v"; python tools/tile_probe.py 467 68 $v 2>&1 | tail -2
done | tee -a $OUT/r02f_variants.log
