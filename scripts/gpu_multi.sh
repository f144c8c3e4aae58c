# GPU box with N GPUs (gpurun --gpus N): the in-process multi-device path on real peers, the frt_main run on N devices,
# and the bench line under torchrun (configs block included: veach / teapot / soup shares per rank)
N=${1:-2}; set -x; mkdir -p gpurun_out
nvidia-smi -L
timeout 900 python -m pytest tests/test_gpu_render.py -m gpu -q -s -x -k "multi_gpu or multi_device" 2>&1 | grep -v "^Loading" > gpurun_out/pytest_multi_n$N.log; tail -5 gpurun_out/pytest_multi_n$N.log
timeout 600 python - > gpurun_out/inprocess_n$N.log 2>&1 <<PY
import sys, time; sys.path.insert(0, '.')
import numpy as np
from first_raytracer_b200 import api
s = api.HostScene("cornell", 16 / 9, 2)
for n in (1, $N):
    s.render(1920, 1080, 8, n_gpus=n)
    t0 = time.perf_counter(); img, st, secs = s.render(1920, 1080, 1024, n_gpus=n); dt = time.perf_counter() - t0
    print(f"renderer<gpu_path_integrator> n_gpus={n}: wall {dt*1e3:.1f} ms, device {st['render_ms']:.1f} ms, {1920*1080*1024/dt/1e6:.0f} Msamples/s e2e, mean {img.mean():.6f}", flush=True)
# ray casting on N devices at once: 2^24 incoherent rays against the teapot, pinned buffers
sc = api.HostScene("teapot", 16 / 9, 2)
rng = np.random.default_rng(1)
info = sc.info
n = 1 << 24
pr, ph = api.Pinned(n, api.RAY_DTYPE), api.Pinned(n, api.HIT_DTYPE)
lo, hi = np.array(info["bounds_lo"]), np.array(info["bounds_hi"])
pr.array["o"] = (lo + rng.random((n, 3)) * (hi - lo)).astype(np.float32)
d = rng.normal(size=(n, 3)); pr.array["d"] = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(np.float32)
for k in (1, $N):
    ctx = api.Context(n=k); ctx.upload(sc)
    ctx.trace(pr.array, hits=ph.array)
    t0 = time.perf_counter(); _, ms = ctx.trace(pr.array, hits=ph.array); dt = time.perf_counter() - t0
    print(f"frt_trace_rays on {k} device(s): device {ms:.2f} ms ({n/ms/1e3:.0f} Mrays/s), e2e {dt*1e3:.1f} ms ({n/dt/1e6:.0f} Mrays/s), hits {int((ph.array['prim_id']>=0).sum())}", flush=True)
    ctx.close()
PY
grep -v Loading gpurun_out/inprocess_n$N.log | grep -E "renderer<|frt_trace" 
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_n$N.log 2> gpurun_out/bench_n${N}_err.log; wc -c gpurun_out/bench_n$N.log; tail -3 gpurun_out/bench_n${N}_err.log
python - <<PY
import json
d = json.loads(open('gpurun_out/bench_n$N.log').read().strip().splitlines()[-1])
print({k: d[k] for k in ('value', 'n_gpus', 'ms_per_step')}, 'e2e', d['e2e']['value'] if d.get('e2e') else None)
for k, v in (d.get('configs') or {}).items():
    if 'casts' in v:
        for c in v['casts']: print('  soup', c['query'], c['order'][:8], round(c['value']), 'Mrays/s', c.get('parity'))
        print('  soup e2e', v.get('e2e'))
    else:
        print(' ', k, round(v['value'], 1), v['unit'], v.get('parity', {}).get('ok'))
PY
