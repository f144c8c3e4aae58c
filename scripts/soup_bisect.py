"""GPU box: find what makes one ray set of the soup cast 60 % slower than another with the same work per ray —
time blocks of the slow set and bisect into the slowest one."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from first_raytracer_b200 import api
dev = torch.device("cuda", 0)
n = 1 << 24
seed = int(sys.argv[1]) if len(sys.argv) > 1 else 1
scene = api.HostScene(soup=api.soup_triangles(10_000_000), fast_bvh=True, max_leaf=4)
ctx = api.Context(devices=[0]); ctx.upload(scene)
g = torch.Generator(device=dev); g.manual_seed(0x5EED0002 + seed)
o = torch.rand((n, 3), device=dev, generator=g)
d = torch.randn((n, 3), device=dev, generator=g); d = d / d.norm(dim=1, keepdim=True)
rays = torch.cat([o, d], 1).contiguous()
hits = torch.empty((n, 4), dtype=torch.float32, device=dev)
torch.cuda.synchronize()

def cast(b, e, reps=2):
    return min(ctx.trace_device(rays.data_ptr() + 24 * b, e - b, hits.data_ptr() + 16 * b) for _ in range(reps))

print("whole set: %.2f ms" % cast(0, n, 3), flush=True)
blk = n // 16
times = [cast(i * blk, (i + 1) * blk) for i in range(16)]
print("16 blocks (ms):", [round(t, 2) for t in times], "sum %.2f" % sum(times), flush=True)
b, e = max(range(16), key=lambda i: times[i]) * blk, 0
e = b + blk
while e - b > 32:
    m = (b + e) // 2
    t0, t1 = cast(b, m), cast(m, e)
    print(f"  [{b},{m}) {t0:.3f} ms   [{m},{e}) {t1:.3f} ms", flush=True)
    b, e = (b, m) if t0 >= t1 else (m, e)
ctx.set_flags(api.FRT_FLAG_COUNTERS)
for i in range(b, e):
    c0 = ctx.counters(); ms = ctx.trace_device(rays.data_ptr() + 24 * i, 1, hits.data_ptr() + 16 * i); c1 = ctx.counters()
    nv = c1["node_visits"] - c0["node_visits"]
    if nv > 500 or ms > 0.2:
        print(json.dumps({"ray": i, "ms": round(ms, 3), "nodes": nv, "tris": c1["tri_tests"] - c0["tri_tests"], "o_d": [float(x) for x in rays[i]],
                          "hit": [float(x) for x in hits[i, :3]] + [int(hits[i, 3].view(torch.int32))]}), flush=True)
# distribution of per-ray node visits on a sample: the heaviest rays of 2^16
ctx.set_flags(api.FRT_FLAG_COUNTERS)
worst = []
for i in range(0, 1 << 12):
    c0 = ctx.counters(); ctx.trace_device(rays.data_ptr() + 24 * i, 1, hits.data_ptr() + 16 * i); c1 = ctx.counters()
    worst.append(c1["node_visits"] - c0["node_visits"])
worst.sort()
print("per-ray node visits over 4096 rays: median", worst[len(worst) // 2], "p99", worst[int(0.99 * len(worst))], "max", worst[-1])
