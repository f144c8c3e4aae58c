# GPU box: A/B of two builds of libfrt.so (the box's copy of the repo is scratch): build/libfrt_exp.so against the shipped one
set -x; mkdir -p gpurun_out
run() {
  SWEEP_LEAF=0 timeout 900 python scripts/wf_sweep.py cornell:256 veach:256 teapot:64 -- "" 2>&1 | grep "^{" | cut -c1-200
  VARIANTS="wide8=0" SOUP_LEAF=3 timeout 600 python scripts/soup_sweep.py 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: continue
    if 'variant' in d: print('soup', d['query'], 'sorted' if d['sorted'] else 'as given', round(d['Mrays_per_s']), 'Mrays/s nodes', round(d['nodes_per_ray'],1), 'tris', round(d['tris_per_ray'],1))
"
}
{ echo "== shipped"; run; cp build/libfrt_exp.so first_raytracer_b200/libfrt.so; echo "== experiment (${EXP_NAME:-build/libfrt_exp.so})"; run; } | tee gpurun_out/ab_lib.log
