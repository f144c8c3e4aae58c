# GPU box, round 2 call D: warp-autonomous wavefront (default) — full parity tests, A/B against the CTA-wide pipeline, any-hit variants on the soup
set -x; mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -s -x 2>&1 | grep -v "^Loading" > gpurun_out/pytest_gpu.log; tail -6 gpurun_out/pytest_gpu.log
timeout 900 python scripts/wf_sweep.py cornell:64 teapot:32 veach:32 -- "wf_mode=0" "wf_mode=1" "wf_mode=1,wf_config=0,wf_slots=512" "wf_mode=1,wf_config=0,wf_slots=768" "wf_mode=1,wf_config=2" "wf_mode=1,wf_config=3,wf_slots=640" "wf_mode=1,wf_config=4,wf_slots=1280" "wf_mode=1,wf_slots=1536" "wf_mode=1,wf_refill=16" "wf_mode=1,wf_refill=28" "wf_mode=1,wf_sort=0" > gpurun_out/wf_sweep_d.log 2>&1; cat gpurun_out/wf_sweep_d.log | cut -c1-330
timeout 300 python bench.py --steps 2 --warmup 3 --no-configs --no-cpu-baseline > gpurun_out/bench_d.log 2> gpurun_out/bench_d_err.log; cut -c1-300 gpurun_out/bench_d.log
VARIANTS="wide8=0;wide8=0,trace_refill_any=24;wide8=0,tune=8,trace_refill_any=24;wide8=0,tune=8,trace_refill_any=28;wide8=0,tune=8,trace_refill_any=24,trace_occ_any=1;wide8=0,tune=10,trace_refill_any=24;wide8=0,tune=8" timeout 1200 python scripts/soup_sweep.py > gpurun_out/soup_sweep_d.log 2> gpurun_out/soup_sweep_d_err.log
python - <<'PY'
import json
for l in open('gpurun_out/soup_sweep_d.log'):
    try: d=json.loads(l)
    except Exception: continue
    if 'variant' in d and d['query']=='shadow': print(f"{d['variant']:60s} {d['query']:8s} sorted={int(d['sorted'])} {d['Mrays_per_s']:7.0f} Mrays/s nodes {d['nodes_per_ray']:.1f} tris {d['tris_per_ray']:.1f} same {d['same_as_first_variant']:.6f}")
PY
tail -3 gpurun_out/soup_sweep_d_err.log
