# GPU box: mesh scenes under the shipped libfrt.so and under build/libfrt_exp.so
set -x; mkdir -p gpurun_out
run() { SWEEP_LEAF=0 timeout 900 python scripts/wf_sweep.py teapot:64 furnace+hdr:64 cornell:64 -- "" 2>&1 | grep "^{" | cut -c1-200; }
{ echo "== shipped"; run; cp build/libfrt_exp.so first_raytracer_b200/libfrt.so; echo "== experiment ${EXP_NAME}"; run; } | tee gpurun_out/ab_mesh.log
