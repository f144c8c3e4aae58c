# GPU box, round 2 call G: launch shape by scene size — parity tests, smoke, pool-size sweep, full bench line + reference arm
set -x; mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -s -x 2>&1 | grep -v "^Loading" > gpurun_out/pytest_gpu.log; tail -5 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | grep -v "^Loading" > gpurun_out/smoke.log; tail -2 gpurun_out/smoke.log
timeout 1500 python scripts/wf_sweep.py cornell:256 veach:256 teapot:64 -- "" "wf_config=1,wf_slots=1280" "wf_config=0,wf_slots=512" "wf_config=0,wf_slots=640" "wf_config=0,wf_slots=384" "wf_config=3,wf_slots=512" "wf_config=1,wf_slots=1024" > gpurun_out/wf_sweep_pool_size.log 2>&1
cat gpurun_out/wf_sweep_pool_size.log | cut -c1-220
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench_err.log; wc -c gpurun_out/bench.log; tail -4 gpurun_out/bench_err.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.log 2> gpurun_out/bench_reference_err.log; tail -c 600 gpurun_out/bench_reference.log
