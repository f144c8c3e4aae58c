# GPU box, round 2 call I: the no-alloc ray cast — its test alone in a fresh process, then after every other GPU test
set -x; mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q -s -x 2>&1 | grep -v "^Loading" > gpurun_out/pytest_no_alloc_alone.log; tail -8 gpurun_out/pytest_no_alloc_alone.log
timeout 2400 python -m pytest tests -m gpu -q -s -x 2>&1 | grep -v "^Loading" > gpurun_out/pytest_gpu.log; tail -8 gpurun_out/pytest_gpu.log
