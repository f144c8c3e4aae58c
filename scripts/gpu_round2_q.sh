# GPU box: the ray-cast knobs again, now that no pathological ray hides the differences
set -x; mkdir -p gpurun_out
VARIANTS="wide8=0;trace_refill=12;trace_refill=16;trace_refill=16,trace_refill_any=20;trace_refill_any=20;trace_refill_any=22;trace_occ_any=1;trace_occ=1,trace_refill=16" timeout 1500 python scripts/soup_sweep.py > gpurun_out/soup_sweep_clean2.log 2> gpurun_out/soup_sweep_clean2_err.log
python - <<'PY'
import json
for l in open('gpurun_out/soup_sweep_clean2.log'):
    try: d=json.loads(l)
    except Exception: continue
    if 'variant' in d: print(f"{d['variant']:44s} {d['query']:8s} sorted={int(d['sorted'])} {d['Mrays_per_s']:7.0f} Mrays/s nodes {d['nodes_per_ray']:.1f} tris {d['tris_per_ray']:.1f} same {d['same_as_first_variant']:.6f}")
PY
tail -3 gpurun_out/soup_sweep_clean2_err.log
