# GPU box, round 2 call C: defaults + sweeps, ncu captures summarised ON the box (the reports stay there: 3 x 50 MB > the 64 MiB that comes back)
set -x; mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -s -x -k "cancel or tonemap or multi_device or counters or pinned" 2>&1 | grep -v "^Loading" > gpurun_out/pytest_c.log; tail -8 gpurun_out/pytest_c.log
timeout 300 python bench.py --steps 2 --warmup 3 --no-configs --no-cpu-baseline > gpurun_out/bench_c.log 2> gpurun_out/bench_c_err.log; cut -c1-900 gpurun_out/bench_c.log
FRT_HOST_TIMING=1 python - > gpurun_out/host_timing.log 2>&1 <<'PY'
import sys; sys.path.insert(0, '.')
from first_raytracer_b200 import api
s = api.HostScene("cornell", 16 / 9, 2)
for spp in (4, 64, 64):
    img, st, secs = s.render(1920, 1080, spp)
    print(f"== spp {spp}: wall {secs * 1e3:.2f} ms, device {st['render_ms']:.2f} ms", file=sys.stderr, flush=True)
PY
grep -v Loading gpurun_out/host_timing.log | tail -24
timeout 900 python scripts/wf_sweep.py cornell:64 teapot:32 veach:32 -- "" "tune=4" "wf_config=0,wf_slots=512" "wf_config=0,wf_slots=512,tune=4" "wf_slots=1152" "wf_slots=1280" "wf_refill=20" "wf_refill=28" > gpurun_out/wf_sweep_c.log 2>&1; cat gpurun_out/wf_sweep_c.log | cut -c1-330
VARIANTS="wide8=0;wide8=0,tune=2;wide8=0,tune=6;wide8=0,tune=6,trace_occ=3,trace_occ_any=3;wide8=0,tune=6,trace_step=2,trace_refill=28;wide8=0,tune=6,trace_occ=1,trace_occ_any=1;wide8=1,tune=6;wide8=1,tune=6,wide8_stride=8" timeout 1200 python scripts/soup_sweep.py > gpurun_out/soup_sweep_c.log 2> gpurun_out/soup_sweep_c_err.log
SOUP_LEAF=2 VARIANTS="wide8=0,tune=6;wide8=1,tune=6" timeout 600 python scripts/soup_sweep.py > gpurun_out/soup_sweep_c_leaf2.log 2>&1
python - <<'PY'
import json
for f in ('gpurun_out/soup_sweep_c.log', 'gpurun_out/soup_sweep_c_leaf2.log'):
    for l in open(f):
        try: d=json.loads(l)
        except Exception: continue
        if 'variant' in d: print(f"{f[-9:-4]} {d['variant']:70s} {d['query']:8s} sorted={int(d['sorted'])} {d['Mrays_per_s']:7.0f} Mrays/s nodes {d['nodes_per_ray']:.1f} tris {d['tris_per_ray']:.1f} same {d['same_as_first_variant']:.6f} frac {d['frac_of_hbm_peak']:.3f}")
PY
tail -3 gpurun_out/soup_sweep_c_err.log
# ncu (one tool per call): each capture only after the same command exited 0 without ncu; summaries as text, reports removed
prof() { # name, kernel regex, command...
  local name=$1 k=$2; shift 2
  "$@" > gpurun_out/prof_${name}_plain.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:$k --launch-skip 1 --launch-count 1 -f -o /tmp/prof_$name "$@" > gpurun_out/prof_${name}_ncu.log 2>&1
  { echo "# ncu --set full --clock-control none --import-source on -k regex:$k (second launch of: $*)"; tail -1 gpurun_out/prof_${name}_plain.log
    python scripts/ncu_summary.py /tmp/prof_$name.ncu-rep; echo; echo "# per source line (scripts/ncu_lines.py)"; python scripts/ncu_lines.py /tmp/prof_$name.ncu-rep 45; } > gpurun_out/ncu_$name.txt 2>&1
  ls -la /tmp/prof_$name.ncu-rep; head -30 gpurun_out/ncu_$name.txt
}
FRT_WIDE8=0 FRT_TUNE=6 prof soup k_trace python scripts/prof_target.py soup 1e7 4194304
FRT_WIDE8=0 FRT_TUNE=6 PROF_ANY=1 prof soup_any k_trace python scripts/prof_target.py soup 1e7 4194304
prof teapot k_wavefront python scripts/prof_target.py teapot 8
prof cornell k_wavefront python scripts/prof_target.py cornell 8
cp /tmp/prof_cornell.ncu-rep gpurun_out/ 2>/dev/null; du -sh gpurun_out
