# GPU box, round 2 call E: hybrid (shared + local) stack in the ray-cast kernels, adaptive stage order in k_wavefront
set -x; mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_trace.py -m gpu -q -s -x 2>&1 | grep -v "^Loading" > gpurun_out/pytest_trace.log; tail -4 gpurun_out/pytest_trace.log
timeout 900 python scripts/wf_sweep.py cornell:64 teapot:32 veach:32 -- "" "wf_slots=1280" "wf_refill=20" > gpurun_out/wf_sweep_e.log 2>&1; cat gpurun_out/wf_sweep_e.log | cut -c1-330
VARIANTS="wide8=0;wide8=0,trace_occ=1;wide8=0,trace_occ=0;wide8=0,tune=8;wide8=0,tune=8,trace_refill_any=24;wide8=0,tune=8,trace_occ_any=1;wide8=0,tune=8,trace_occ_any=0,trace_refill_any=24;wide8=0,tune=10;wide8=1,tune=8;wide8=1,tune=8,wide8_stride=8,trace_occ=1" timeout 1200 python scripts/soup_sweep.py > gpurun_out/soup_sweep_e.log 2> gpurun_out/soup_sweep_e_err.log
python - <<'PY'
import json
for l in open('gpurun_out/soup_sweep_e.log'):
    try: d=json.loads(l)
    except Exception: continue
    if 'variant' in d: print(f"{d['variant']:60s} {d['query']:8s} sorted={int(d['sorted'])} {d['Mrays_per_s']:7.0f} Mrays/s nodes {d['nodes_per_ray']:.1f} tris {d['tris_per_ray']:.1f} same {d['same_as_first_variant']:.6f}")
PY
tail -3 gpurun_out/soup_sweep_e_err.log
FRT_WIDE8=0 python scripts/prof_target.py soup 1e7 4194304 > gpurun_out/prof_soup_plain.log 2>&1 &&
FRT_WIDE8=0 ncu --set full --clock-control none --import-source on -k regex:k_trace --launch-skip 1 --launch-count 1 -f -o /tmp/prof_soup python scripts/prof_target.py soup 1e7 4194304 > gpurun_out/prof_soup_ncu.log 2>&1
{ tail -1 gpurun_out/prof_soup_plain.log; python scripts/ncu_summary.py /tmp/prof_soup.ncu-rep; python scripts/ncu_lines.py /tmp/prof_soup.ncu-rep 30; } > gpurun_out/ncu_soup_e.txt 2>&1; head -24 gpurun_out/ncu_soup_e.txt
