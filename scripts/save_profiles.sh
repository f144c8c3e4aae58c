# copy the last gpu_round.sh results from gpurun_out/ (scratch) into profiles/ (tracked), round prefix $1 (default r01)
R=${1:-r01}
cp gpurun_out/pytest_gpu.log profiles/${R}_pytest_gpu.log; cp gpurun_out/smoke.log profiles/${R}_smoke.log
tail -1 gpurun_out/bench.log > profiles/${R}_bench_n1.json
cp gpurun_out/launches.csv profiles/${R}_ncu_launches.csv
[ -s gpurun_out/scenes.log ] && cp gpurun_out/scenes.log profiles/${R}_scenes.log
{ echo "# ncu --set full --clock-control none --import-source on, k_wavefront, bench.py --spp 8 (one launch = whole 1080p x 8 spp tile)"
  python scripts/ncu_summary.py gpurun_out/prof.ncu-rep; echo; echo "# per source line (scripts/ncu_lines.py)"
  python scripts/ncu_lines.py gpurun_out/prof.ncu-rep 40; } > profiles/${R}_ncu_wavefront.txt 2>/dev/null
