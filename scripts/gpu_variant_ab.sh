#!/bin/bash
# A/B of tuning builds (libqmmx_mc_<tag>.so beside the product library) on C2 / C3 / C2-HL, one B200 (run under gpurun)
#   bash scripts/gpu_variant_ab.sh <tag> [<tag> ...]      ("base" = the product library)
mkdir -p gpurun_out
for tag in "$@"; do
  lib=$PWD/qmmx_monolithic_monte_carlo_b200/libqmmx_mc_$tag.so
  [ "$tag" = base ] && lib=$PWD/qmmx_monolithic_monte_carlo_b200/libqmmx_mc.so
  for w in ${WL:-c2 c3 c2hl}; do
    QMMX_LIB=$lib timeout 300 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --no-sub > gpurun_out/ab_${tag}_$w.json 2> gpurun_out/ab_${tag}_$w.err || tail -3 gpurun_out/ab_${tag}_$w.err
    python -c "
import json
d = json.loads(open('gpurun_out/ab_${tag}_$w.json').read().strip().splitlines()[-1])
print('$tag $w ms_per_step %.4f value %.4g' % (d['ms_per_step'], d['value']), (d.get('result_check') or {}).get('n_trades'))"
  done
done
