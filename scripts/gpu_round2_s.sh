# GPU box: dynamic-fetch threshold and pool size of the mesh-scene shape (four 128-thread CTAs, 4 triangles per leaf)
set -x; mkdir -p gpurun_out
SWEEP_LEAF=0 timeout 1200 python scripts/wf_sweep.py teapot:64 -- "" "wf_refill=12" "wf_refill=16" "wf_refill=24" "wf_refill=28" "wf_slots=224" "wf_slots=192" "wf_slots=288,wf_config=4" "wf_sort=0" "wf_sort=3" > gpurun_out/wf_sweep_teapot.log 2>&1
grep "^{" gpurun_out/wf_sweep_teapot.log | cut -c1-210
