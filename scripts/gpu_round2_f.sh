# GPU box, round 2 call F: cleaned-up kernels — full parity tests, smoke, any-hit occupancy, full bench line
set -x; mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -s -x 2>&1 | grep -v "^Loading" > gpurun_out/pytest_gpu.log; tail -5 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | grep -v "^Loading" > gpurun_out/smoke.log; tail -2 gpurun_out/smoke.log
VARIANTS="wide8=0;wide8=0,trace_occ_any=0;wide8=0,trace_occ_any=1;wide8=0,trace_occ_any=0,trace_refill_any=24;wide8=0,trace_occ_any=1,trace_refill_any=24;wide8=0,trace_occ=1" timeout 1200 python scripts/soup_sweep.py > gpurun_out/soup_sweep_f.log 2> gpurun_out/soup_sweep_f_err.log
python - <<'PY'
import json
for l in open('gpurun_out/soup_sweep_f.log'):
    try: d=json.loads(l)
    except Exception: continue
    if 'variant' in d: print(f"{d['variant']:60s} {d['query']:8s} sorted={int(d['sorted'])} {d['Mrays_per_s']:7.0f} Mrays/s nodes {d['nodes_per_ray']:.1f} tris {d['tris_per_ray']:.1f} same {d['same_as_first_variant']:.6f}")
PY
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench.log 2> gpurun_out/bench_err.log; wc -c gpurun_out/bench.log; tail -4 gpurun_out/bench_err.log
