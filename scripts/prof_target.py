"""GPU box: a minimal run of one kernel for ncu (two launches: the first warms up, profile the second).
  python scripts/prof_target.py soup [n_tri] [n_rays] | teapot [spp] | cornell [spp] | veach [spp]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: F401
from first_raytracer_b200 import api
what = sys.argv[1]
if what == "soup":
    import torch
    n_tri = int(float(sys.argv[2])) if len(sys.argv) > 2 else 10_000_000
    n = int(sys.argv[3]) if len(sys.argv) > 3 else 1 << 22
    scene = api.HostScene(soup=api.soup_triangles(n_tri), fast_bvh=True, max_leaf=int(os.environ.get("SOUP_LEAF", "3")))
    ctx = api.Context(devices=[0]); ctx.upload(scene)
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev); g.manual_seed(0x5EED0002)
    o = torch.rand((n, 3), device=dev, generator=g)
    d = torch.randn((n, 3), device=dev, generator=g); d = d / d.norm(dim=1, keepdim=True)
    rays = torch.cat([o, d], 1).contiguous()
    hits = torch.empty((n, 4), dtype=torch.float32, device=dev)
    any_hit = os.environ.get("PROF_ANY", "0") == "1"
    for _ in range(2):
        ms = ctx.trace_device(rays.data_ptr(), n, hits.data_ptr(), any_hit)
    print(f"soup {n_tri} tris, {n} rays: {n / ms / 1e3:.0f} Mrays/s")
else:
    spp = int(sys.argv[2]) if len(sys.argv) > 2 else 8
    scene = api.HostScene(what, 16 / 9, 0)  # leaf size by scene size, as the bench
    ctx = api.Context(devices=[0]); ctx.upload(scene)
    for _ in range(2):
        ctx.clear_film(); ctx.render_tile(1920, 1080, 0, spp, seed=1)
        _, st = ctx.readback(want_image=False)
    print(f"{what} 1080p x {spp} spp: {1920 * 1080 * spp / st['render_ms'] / 1e3:.0f} Msamples/s, stages {[round(st[k + '_ms'], 1) for k in ('generate', 'extend', 'shade', 'connect')]}")
ctx.close()
