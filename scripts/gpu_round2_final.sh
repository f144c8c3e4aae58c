# GPU box: the round's closing evidence on one GPU — parity tests, smoke, the full bench line, the reference arm
set -x; mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -s -x 2>&1 | grep -v "^Loading" > gpurun_out/pytest_gpu.log; tail -6 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | grep -v "^Loading" > gpurun_out/smoke.log; tail -2 gpurun_out/smoke.log
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench_err.log; wc -c gpurun_out/bench.log; tail -4 gpurun_out/bench_err.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.log 2> gpurun_out/bench_reference_err.log; tail -c 300 gpurun_out/bench_reference.log
