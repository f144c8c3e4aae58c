# GPU box: soup casts under several builds of libfrt.so: the shipped one, then every build/libfrt_exp*.so
set -x; mkdir -p gpurun_out
run() {
  VARIANTS="wide8=0" SOUP_LEAF=3 timeout 600 python scripts/soup_sweep.py 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: continue
    if 'variant' in d: print('soup', d['query'], 'sorted' if d['sorted'] else 'as given', round(d['Mrays_per_s']), 'Mrays/s nodes', round(d['nodes_per_ray'],1), 'tris', round(d['tris_per_ray'],1))
"
}
{ echo "== shipped"; run; for f in build/libfrt_exp*.so; do cp $f first_raytracer_b200/libfrt.so; echo "== $f"; run; done; } | tee gpurun_out/ab_soup.log
