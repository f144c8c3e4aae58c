# GPU box: four 128-thread CTAs per SM for the path kernel (config 4) against the current shapes
set -x; mkdir -p gpurun_out
timeout 1500 python scripts/wf_sweep.py cornell:256 veach:256 teapot:64 -- "" "wf_config=4,wf_slots=320" "wf_config=4,wf_slots=256" "wf_config=4,wf_slots=288" "wf_config=0,wf_slots=512" > gpurun_out/wf_sweep_128x4.log 2>&1
python - <<'PY'
import json
for l in open('gpurun_out/wf_sweep_128x4.log'):
    try: d = json.loads(l)
    except Exception: print(l[:200]); continue
    print(f"{d['scene']:8s} {d['env']:28s} {d.get('Msamples_per_s')} stages {d.get('stage_ms')} iters {d.get('iters')}")
PY
