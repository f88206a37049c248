#!/bin/bash
# bench.py at N = 8 and 4 on one box (run under `gpurun --gpus 8`)
mkdir -p gpurun_out
for n in 8 4; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29800 + n)) \
      bench.py --gpus $n --steps 20 --warmup 3 > gpurun_out/bench_n${n}_full_line.json 2> gpurun_out/bench_n$n.err || tail -5 gpurun_out/bench_n$n.err
  python - <<PY
import json
try:
    d = json.loads(open("gpurun_out/bench_n${n}_full_line.json").read().strip().splitlines()[-1])
    print("N=$n value %.4g ms %.4f e2e %.4g (%.1f ms, probe %.1f GB/s aggregate, e2e/probe %.3f) c3 %.4g %.2f ms digest %s c4 %.4g c5f %.4g" % (
        d["value"], d["ms_per_step"], d["e2e"]["value"], d["e2e"]["ms_per_step"], d["e2e"]["h2d_probe"]["gb_per_s_aggregate"],
        d["e2e"]["h2d_probe"]["e2e_over_probe"], d["c3"]["value"], d["c3"]["ms"], d["c3"]["agg_digest"], d["c4"]["value"], d["c5f"]["value"]))
except Exception as e:
    print("N=$n no line:", e)
PY
done
