# GPU box: layout of the sort key (bits per axis of the origin cell : of the direction) on the soup
set -x; mkdir -p gpurun_out
SOUP_LEAF=3 VARIANTS="wide8=0;sort_bits=9:1;sort_bits=8:2;sort_bits=7:3;sort_bits=6:4;sort_bits=5:5;sort_bits=7:2;sort_bits=6:3;sort_bits=5:3;sort_bits=4:4" timeout 1500 python scripts/soup_sweep.py 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    try: d = json.loads(l)
    except Exception: continue
    if 'variant' in d and d['sorted']: print(d['variant'], d['query'], 'sorted', round(d['Mrays_per_s']), 'Mrays/s same', d['same_as_first_variant'])
" | tee gpurun_out/soup_sort_bits.log
