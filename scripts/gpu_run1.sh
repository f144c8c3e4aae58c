set -x; mkdir -p gpurun_out
nvidia-smi -L | head -2
python -m pytest tests -m gpu -x -q 2>&1 | tail -40 > gpurun_out/r1_pytest.log; tail -15 gpurun_out/r1_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r1_smoke.log 2>&1; tail -5 gpurun_out/r1_smoke.log
timeout 600 python bench.py --steps 1 --warmup 1 --spp 64 > gpurun_out/r1_bench64.log 2>&1; tail -3 gpurun_out/r1_bench64.log
timeout 900 python bench.py --steps 2 --warmup 3 > gpurun_out/r1_bench.log 2>&1; tail -3 gpurun_out/r1_bench.log
