#!/bin/bash
# one kernel iteration on the GPU box: tile tests, C2/C3 timings, ncu summary + SASS counters of the tile scan
tag=${1:-it}
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q -k "tile" 2>&1 | tail -4
for w in c2 c3; do
  timeout 300 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline --no-e2e > gpurun_out/${tag}_${w}.json 2> gpurun_out/${tag}_${w}.err || tail -5 gpurun_out/${tag}_${w}.err
  python -c "
import json
d = json.loads(open('gpurun_out/${tag}_${w}.json').read().strip().splitlines()[-1])
print('$w ms_per_step %.4f value %.4g' % (d['ms_per_step'], d['value']), d.get('result_check'))"
done
bash profiles/capture.sh $tag scan:tile_scan 2>&1 | tail -3
QMMX_NO_TMA=1 timeout 300 python bench.py --workload c2 --steps 10 --warmup 3 --no-cpu-baseline --no-e2e 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(\"c2 (cp.async staging) ms_per_step %.4f\" % d[\"ms_per_step\"])"
grep -E "^(smsp__inst_executed.sum|smsp__issue_active|smsp__thread_inst|gpu__time)" gpurun_out/${tag}_scan.summary.txt
