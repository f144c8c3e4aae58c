# GPU box: DRAM bytes of the full-size k_wavefront launch (the one bench.py times) — one ncu pass of three metrics,
# after the same command exited 0 without ncu.  Output: gpurun_out/traffic_1024.csv (bench.py's roofline.traffic)
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --spp ${SPP:-1024} --no-e2e --no-cpu-baseline --no-roofline"
$CMD > gpurun_out/traffic_plain.log 2>&1 &&
timeout 300 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none \
  -k regex:k_wavefront --launch-skip 1 --launch-count 1 --csv --log-file gpurun_out/traffic_1024.csv $CMD > gpurun_out/traffic_ncu.log 2>&1
tail -5 gpurun_out/traffic_1024.csv
