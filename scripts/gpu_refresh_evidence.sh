mkdir -p gpurun_out
bash profiles/capture.sh r2 scan:tile_scan policy:tile_scan fused:tile_fused scanhl:lean_scan sweep:lean_sweep boot:lean_fused_boot retrain:lean_scan 2>&1 | tee gpurun_out/capture_refresh.log | tail -30
timeout 300 python bench.py --steps 20 --warmup 3 --no-sub --no-cpu-baseline > gpurun_out/l_plain.json 2> gpurun_out/l_plain.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c2_bench.csv python bench.py --steps 20 --warmup 3 --no-sub --no-cpu-baseline --no-e2e > gpurun_out/l_ncu.log 2>&1
tail -3 gpurun_out/launches_c2_bench.csv
