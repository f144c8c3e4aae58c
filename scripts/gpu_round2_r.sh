# GPU box: triangles per leaf of the soup's tree (fast binned builder), clean ray sets
set -x; mkdir -p gpurun_out
for leaf in 2 3 4 6 8; do SOUP_LEAF=$leaf VARIANTS="wide8=0" timeout 600 python scripts/soup_sweep.py 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: continue
    if 'variant' in d: print('leaf $leaf', d['query'], 'sorted' if d['sorted'] else 'as given', round(d['Mrays_per_s']), 'Mrays/s nodes', round(d['nodes_per_ray'],1), 'tris', round(d['tris_per_ray'],1))
    elif 'soup' in d: print('leaf $leaf', 'build_s', round(d['build_s'],1), 'nodes', d['wide4_nodes'], 'depth', d['depth4'])
"; done | tee gpurun_out/soup_leaf.log
