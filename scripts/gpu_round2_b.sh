# GPU box, round 2 call B: cancel fix check, kernel variant sweeps, ncu captures of k_trace (soup) and k_wavefront (teapot, cornell)
set -x; mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -s -x -k "cancel or tonemap or multi_device or wide8 or counters or pinned" 2>&1 | grep -v "^Loading" > gpurun_out/pytest_b.log; tail -12 gpurun_out/pytest_b.log
timeout 300 python bench.py --steps 2 --warmup 3 --no-configs --no-cpu-baseline > gpurun_out/bench_b.log 2> gpurun_out/bench_b_err.log; cut -c1-700 gpurun_out/bench_b.log
timeout 900 python scripts/wf_sweep.py cornell:64 teapot:32 veach:32 -- "" "wf_config=14" "wf_config=15" "wf_config=16" "wf_config=17" "wf_config=18" "wf_config=0,wf_slots=512" "wf_config=19,wf_slots=640" "wf_config=20,wf_slots=768" "tune=4" > gpurun_out/wf_sweep.log 2>&1; cat gpurun_out/wf_sweep.log | cut -c1-330
VARIANTS="wide8=0;wide8=0,trace_refill=31;wide8=0,trace_step=1;wide8=0,trace_step=1,trace_refill=31;wide8=0,trace_step=2,trace_refill=28;wide8=0,tune=2;wide8=0,tune=3;wide8=0,trace_occ=1;wide8=0,trace_occ=2;wide8=0,trace_step=1,trace_refill=31,trace_occ=1,tune=2;wide8=0,trace_refill_any=24;wide8=0,trace_refill_any=31,trace_step=1;wide8=0,trace_occ_any=1,trace_refill_any=31,trace_step=1;wide8=1,wide8_stride=8;wide8=1,wide8_stride=8,trace_step=1,trace_refill=31;wide8=1,trace_step=1,trace_refill=31;wide8=1,wide8_stride=8,trace_step=1,trace_refill=31,trace_refill_any=31,tune=2" timeout 1200 python scripts/soup_sweep.py > gpurun_out/soup_sweep_b.log 2> gpurun_out/soup_sweep_b_err.log; python - <<'PY'
import json
for l in open('gpurun_out/soup_sweep_b.log'):
    d=json.loads(l)
    if 'variant' in d: print(f"{d['variant']:90s} {d['query']:8s} sorted={int(d['sorted'])} {d['Mrays_per_s']:7.0f} Mrays/s nodes {d['nodes_per_ray']:.1f} tris {d['tris_per_ray']:.1f} same {d['same_as_first_variant']:.6f}")
PY
tail -3 gpurun_out/soup_sweep_b_err.log
SOUP_LEAF=2 VARIANTS="wide8=0;wide8=0,trace_step=1,trace_refill=31;wide8=1,wide8_stride=8,trace_step=1,trace_refill=31" timeout 600 python scripts/soup_sweep.py > gpurun_out/soup_sweep_leaf2.log 2>&1; python - <<'PY'
import json
for l in open('gpurun_out/soup_sweep_leaf2.log'):
    try: d=json.loads(l)
    except Exception: continue
    if 'variant' in d: print(f"leaf2 {d['variant']:80s} {d['query']:8s} sorted={int(d['sorted'])} {d['Mrays_per_s']:7.0f} Mrays/s nodes {d['nodes_per_ray']:.1f} tris {d['tris_per_ray']:.1f}")
PY
# ncu (one tool per call): each capture only after the same command exited 0 without ncu
FRT_WIDE8=0 python scripts/prof_target.py soup 1e7 4194304 > gpurun_out/prof_soup_plain.log 2>&1 &&
FRT_WIDE8=0 ncu --set full --clock-control none --import-source on -k regex:k_trace --launch-skip 1 --launch-count 1 -f -o gpurun_out/prof_soup python scripts/prof_target.py soup 1e7 4194304 > gpurun_out/prof_soup_ncu.log 2>&1
python scripts/prof_target.py teapot 8 > gpurun_out/prof_teapot_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_wavefront --launch-skip 1 --launch-count 1 -f -o gpurun_out/prof_teapot python scripts/prof_target.py teapot 8 > gpurun_out/prof_teapot_ncu.log 2>&1
python scripts/prof_target.py cornell 8 > gpurun_out/prof_cornell_plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_wavefront --launch-skip 1 --launch-count 1 -f -o gpurun_out/prof_cornell python scripts/prof_target.py cornell 8 > gpurun_out/prof_cornell_ncu.log 2>&1
tail -2 gpurun_out/prof_*_plain.log gpurun_out/prof_*_ncu.log; ls -la gpurun_out | tail -12
