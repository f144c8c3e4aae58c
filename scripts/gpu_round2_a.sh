# GPU box, round 2 call A: parity tests, smoke, soup sweep, a short bench
set -x; mkdir -p gpurun_out
nvidia-smi -L | head -2; nproc
timeout 2400 python -m pytest tests -m gpu -q -s -x 2>&1 | grep -v "^Loading" > gpurun_out/pytest_gpu.log; tail -30 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | grep -v "^Loading" > gpurun_out/smoke.log; tail -3 gpurun_out/smoke.log
timeout 900 python scripts/soup_sweep.py > gpurun_out/soup_sweep.log 2> gpurun_out/soup_sweep_err.log; cat gpurun_out/soup_sweep.log | cut -c1-420; tail -3 gpurun_out/soup_sweep_err.log
timeout 900 python bench.py --steps 2 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench_err.log; wc -c gpurun_out/bench.log; cut -c1-3000 gpurun_out/bench.log; tail -5 gpurun_out/bench_err.log
