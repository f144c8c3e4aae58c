"""GPU box: the soup cast over ray sets that differ only in their seed (why do some sets run 60 % slower?)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from first_raytracer_b200 import api
dev = torch.device("cuda", 0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 24
scene = api.HostScene(soup=api.soup_triangles(10_000_000), fast_bvh=True, max_leaf=4)
ctx = api.Context(devices=[0]); ctx.upload(scene)
hits = torch.empty((n, 4), dtype=torch.float32, device=dev)
for seed in range(6):
    for gen in ("torch", "torch_sep"):
        g = torch.Generator(device=dev); g.manual_seed(0x5EED0002 + seed)
        o = torch.rand((n, 3), device=dev, generator=g)
        if gen == "torch_sep":  # directions from a second, unrelated generator
            g = torch.Generator(device=dev); g.manual_seed(977 * seed + 13)
        d = torch.randn((n, 3), device=dev, generator=g); d = d / d.norm(dim=1, keepdim=True)
        rays = torch.cat([o, d], 1).contiguous()
        ctx.set_flags(0)
        ms = min(ctx.trace_device(rays.data_ptr(), n, hits.data_ptr()) for _ in range(3))
        ms_s = min(ctx.trace_device(rays.data_ptr(), n, hits.data_ptr(), sort=True) for _ in range(2))
        ctx.set_flags(api.FRT_FLAG_COUNTERS)
        c0 = ctx.counters(); ctx.trace_device(rays.data_ptr(), n, hits.data_ptr()); c1 = ctx.counters()
        ctx.set_flags(0)
        h = hits[:, 3].view(torch.int32)
        print(json.dumps({"seed": seed, "gen": gen, "n": n, "Mrays": round(n / ms / 1e3), "Mrays_sorted": round(n / ms_s / 1e3),
                          "nodes_per_ray": round((c1["node_visits"] - c0["node_visits"]) / n, 2), "tris_per_ray": round((c1["tri_tests"] - c0["tri_tests"]) / n, 2),
                          "hit_frac": round(float((h >= 0).float().mean()), 4), "mean_t": round(float(hits[:, 0][h >= 0].mean()), 4),
                          "o_mean": [round(float(x), 4) for x in o.mean(0)], "d_mean": [round(float(x), 4) for x in d.mean(0)],
                          "nan": int(torch.isnan(rays).sum()), "ptr_mod_4k": rays.data_ptr() % 4096}), flush=True)
        del o, d, rays
