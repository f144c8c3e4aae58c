# GPU box: parity tests, smoke, bench, then ncu launch list + full capture (B200_PROFILING.md order:
# every ncu pass only after the same command exited 0 without ncu)
set -x; mkdir -p gpurun_out
nvidia-smi -L | head -2
python -m pytest tests -m gpu -q -s 2>&1 | grep -v "^Loading" > gpurun_out/pytest_gpu.log; tail -5 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | grep -v "^Loading" > gpurun_out/smoke.log; tail -3 gpurun_out/smoke.log
timeout 900 python bench.py --steps 2 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench_err.log; wc -l gpurun_out/bench.log; cut -c1-1500 gpurun_out/bench.log
SPP=${SPP:-8} bash scripts/gpu_profile.sh
