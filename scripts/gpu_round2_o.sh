# GPU box: leaf size of the reference-shaped tree on the mesh scene, with the 128x4 shape as default
set -x; mkdir -p gpurun_out
for leaf in 1 2 4 8; do SWEEP_LEAF=$leaf timeout 600 python scripts/wf_sweep.py teapot:64 cornell:64 -- "" ; done > gpurun_out/wf_sweep_leaf.log 2>&1
grep "^{" gpurun_out/wf_sweep_leaf.log | cut -c1-200
timeout 600 python -m pytest tests/test_gpu_render.py -m gpu -q -x 2>&1 | tail -3
