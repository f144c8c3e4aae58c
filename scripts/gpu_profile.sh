# GPU box: ncu launch list + one full capture of the persistent wavefront kernel (B200_PROFILING.md:
# each ncu pass only after the same command exited 0 without ncu)
# Keep SPP small (8): --set full replays the launch ~40 times with instrumented passes (a 1024-spp launch takes ~19 min
# and two copies of the report exceed gpurun_out's 64 MiB); DRAM bytes of the full-size launch: scripts/gpu_traffic.sh
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --spp ${SPP:-8} --no-e2e --no-cpu-baseline --no-roofline"
$CMD > gpurun_out/prof_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/prof_ncu1.log 2>&1
$CMD > gpurun_out/prof_plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:${KERNEL:-k_wavefront} --launch-skip ${SKIP:-1} --launch-count ${COUNT:-1} -f -o gpurun_out/prof $CMD > gpurun_out/prof_ncu2.log 2>&1
tail -1 gpurun_out/prof_plain.log | cut -c1-200; tail -3 gpurun_out/prof_ncu2.log; ls -la gpurun_out | tail -8
