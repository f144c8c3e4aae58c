# GPU box: one mixed trace pass (FRT_WF_MIXED=1) against the two back-to-back passes
set -x; mkdir -p gpurun_out
SWEEP_LEAF=0 timeout 1500 python scripts/wf_sweep.py cornell:256 veach:256 teapot:64 -- "" "wf_mixed=1" "wf_mixed=1,wf_refill=16" "wf_mixed=1,wf_refill=24" > gpurun_out/wf_sweep_mixed.log 2>&1
python - <<'PY'
import json
for l in open('gpurun_out/wf_sweep_mixed.log'):
    try: d = json.loads(l)
    except Exception: print(l[:200]); continue
    print(f"{d['scene']:8s} {d['env']:28s} {d.get('Msamples_per_s')} stages {d.get('stage_ms')} iters {d.get('iters')} mean {d.get('mean')}")
PY
FRT_WF_MIXED=1 timeout 900 python -m pytest tests/test_gpu_render.py -m gpu -q -x 2>&1 | tail -3
