#!/bin/bash
# compute-sanitizer over every kernel family at small sizes (run under gpurun; ONE tool per call)
tool=${1:-memcheck}; P=${2:-4099}
mkdir -p gpurun_out
log=gpurun_out/sanitizer_${tool}.log
: > $log
for c in scan scanhl policy sweep fused boot genboot gengbm jitter lastbars regime retrain; do
  echo "=== prof_case $c $P" >> $log
  timeout 900 compute-sanitizer --tool $tool --error-exitcode 7 python profiles/prof_case.py $c $P >> $log 2>&1; echo "exit $?" >> $log
done
echo "=== pytest subset" >> $log
timeout 1500 compute-sanitizer --tool $tool --error-exitcode 7 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "tile_kernel_equals or edge_cases or reference_golden or tile_fused or hl_exit" >> $log 2>&1; echo "exit $?" >> $log
grep -E "^=== |^exit |ERROR SUMMARY|passed|failed|Error|RACECHECK SUMMARY" $log | head -80
