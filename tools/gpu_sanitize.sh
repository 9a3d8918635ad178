#!/bin/bash
# compute-sanitizer over the smoke test, one small frame sequence per algorithm, the exact variants and a 3-band group run
# (tools/sanitize_frames.py). memcheck: out-of-bounds / misaligned global, shared and local accesses, leaks of device
# allocations; racecheck: shared-memory hazards (the TMA-filled light stages and the running totals of k_eval_seg);
# synccheck: barrier misuse. Logs go to gpurun_out/<tag>_sanitize_<tool>.log; copy them to profiles/.
TAG=${1:-r02}
OUT=gpurun_out
mkdir -p $OUT
for tool in memcheck racecheck synccheck; do
  extra=""
  [ $tool = memcheck ] && extra="--leak-check full"
  timeout 1500 compute-sanitizer --tool $tool $extra --error-exitcode 9 --print-limit 20 python tools/sanitize_frames.py > $OUT/${TAG}_sanitize_$tool.log 2>&1
  echo "$tool rc=$?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|LEAK SUMMARY|sanitize: all ok" $OUT/${TAG}_sanitize_$tool.log
done
