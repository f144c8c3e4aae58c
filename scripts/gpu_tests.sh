# GPU box: parity tests + smoke (run with: gpurun -- 'bash scripts/gpu_tests.sh')
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -s 2>&1 | grep -v "^Loading" > gpurun_out/pytest_gpu.log; tail -45 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | grep -v "^Loading" > gpurun_out/smoke.log; tail -4 gpurun_out/smoke.log
