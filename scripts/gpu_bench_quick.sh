mkdir -p gpurun_out
bash scripts/gpu_tests.sh
timeout 900 python bench.py --steps 2 --warmup 3 > gpurun_out/bench.log 2>&1; tail -1 gpurun_out/bench.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print({k:d[k] for k in ('value','mrays_per_s','ms_per_step')}, d['roofline']['kernel_ms'], d['e2e']['value'])"
