#!/bin/bash
# A/B of the tile-event kernels against the lockstep lean kernels on one B200 (run under gpurun).
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm --format=csv
echo "== tests (tile + neighbours)"
timeout 900 python -m pytest tests -m gpu -x -q -k "tile or random_paths or lean_kernels or lr_gates or edge or fused_equals" 2>&1 | tail -15
echo "== bench A/B"
for w in c2 c3; do
  for nt in 1 0; do
    echo "-- workload $w QMMX_NO_TILE=$nt"
    QMMX_NO_TILE=$nt timeout 600 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/ab_${w}_notile${nt}.json 2> gpurun_out/ab_${w}_notile${nt}.err || tail -5 gpurun_out/ab_${w}_notile${nt}.err
    python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/ab_${w}_notile${nt}.json").read().strip().splitlines()[-1])
    print("ms_per_step", d["ms_per_step"], "value %.4g" % d["value"], "frac", d["roofline"]["frac"], d.get("result_check"))
except Exception as e:
    print("no line:", e)
PY
  done
done
