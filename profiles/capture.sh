#!/bin/bash
# ncu captures of every kernel of libqmmx_mc.so (run under gpurun, one GPU).  Each case runs plain first (must exit 0),
# then once under `ncu --set full` for the kernel named by the regex, then the summary is written next to the report.
#   bash profiles/capture.sh <tag> case:regex[:paths] ...
set -u
tag=$1; shift
mkdir -p gpurun_out
PIPES=sm__inst_executed_pipe_alu.sum,sm__inst_executed_pipe_fma.sum,sm__inst_executed_pipe_fmaheavy.sum,sm__inst_executed_pipe_fp64.sum,sm__inst_executed_pipe_lsu.sum,sm__inst_executed_pipe_xu.sum,sm__inst_executed_pipe_cbu.sum,sm__inst_executed_pipe_adu.sum,sm__inst_executed_pipe_uniform.sum,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio
for spec in "$@"; do
  IFS=: read -r case regex paths <<< "$spec"
  out=gpurun_out/${tag}_${case}
  python profiles/prof_case.py "$case" ${paths:-} > "$out.plain.log" 2>&1 || { echo "PLAIN RUN FAILED: $case"; tail -5 "$out.plain.log"; continue; }
  cat "$out.plain.log"
  timeout 600 ncu --set full --metrics $PIPES --clock-control none --import-source on -k "regex:$regex" -s 2 -c 1 -f -o "$out" \
      python profiles/prof_case.py "$case" ${paths:-} > "$out.ncu.log" 2>&1 || { echo "NCU FAILED: $case"; tail -5 "$out.ncu.log"; continue; }
  python profiles/ncu_summary.py "$out.ncu-rep" > "$out.summary.txt" 2>&1
  head -3 "$out.summary.txt"
  # per-SASS-instruction counters (needs -lineinfo): kept compressed; the report itself only when it is small
  ncu -i "$out.ncu-rep" --page source --csv 2>/dev/null | gzip -9 > "$out.source.csv.gz"
  if [ "$(stat -c %s "$out.ncu-rep")" -gt 12000000 ]; then rm -f "$out.ncu-rep"; fi
done
