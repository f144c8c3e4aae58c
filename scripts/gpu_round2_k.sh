# 2-GPU box: soup cast per rank alone vs together (why config 5 scales below N), pipelined frt_trace_rays e2e, N=2 bench
set -x; mkdir -p gpurun_out
timeout 600 python scripts/soup_ranks.py 2 > gpurun_out/soup_ranks_n2.log 2>&1; cat gpurun_out/soup_ranks_n2.log | cut -c1-400
timeout 600 python -m pytest tests/test_gpu_trace.py -m gpu -q -x 2>&1 | tail -3
for c in 1048576 2097152 4194304 1000000000; do FRT_TRACE_CHUNK=$c timeout 300 python - <<PY
import sys, time; sys.path.insert(0, '.')
import numpy as np
from first_raytracer_b200 import api
sc = api.HostScene(soup=api.soup_triangles(10_000_000), fast_bvh=True, max_leaf=4)
rng = np.random.default_rng(1); n = 1 << 24
pr, ph = api.Pinned(n, api.RAY_DTYPE), api.Pinned(n, api.HIT_DTYPE)
pr.array["o"] = rng.random((n, 3)).astype(np.float32)
d = rng.normal(size=(n, 3)); pr.array["d"] = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
for k in (1, 2):
    ctx = api.Context(n=k); ctx.upload(sc)
    ctx.trace(pr.array, hits=ph.array)
    best = 1e9
    for _ in range(3):
        t0 = time.perf_counter(); _, ms = ctx.trace(pr.array, hits=ph.array); best = min(best, time.perf_counter() - t0)
    print(f"chunk $c: frt_trace_rays soup on {k} device(s): device {ms:.2f} ms, e2e {best*1e3:.1f} ms ({n/best/1e6:.0f} Mrays/s), hits {int((ph.array['prim_id']>=0).sum())}", flush=True)
    ctx.close()
PY
done 2>&1 | grep -v Loading | tee gpurun_out/trace_pipeline_n2.log
bash scripts/gpu_multi.sh 2
