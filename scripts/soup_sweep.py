"""GPU box: config 5 (10^7-triangle soup) ray-cast throughput over the kernel variants, one JSON line each.
env SOUP_TRIS (1e7), SOUP_RAYS (2**24), SOUP_LEAF (4), VARIANTS ("wide8=0;wide8=1,prefetch=0;wide8=1,prefetch=1")"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from first_raytracer_b200 import api

n_tri = int(float(os.environ.get("SOUP_TRIS", "1e7")))
n_rays = int(os.environ.get("SOUP_RAYS", str(2 ** 24)))
leaf = int(os.environ.get("SOUP_LEAF", "4"))
variants = os.environ.get("VARIANTS", "wide8=0;wide8=1,prefetch=0;wide8=1,prefetch=1").split(";")
PEAK = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json"))).get("hbm_gbs", 6650.0)

t0 = time.time(); tri = api.soup_triangles(n_tri); t1 = time.time()
scene = api.HostScene(soup=tri, fast_bvh=True, max_leaf=leaf); t2 = time.time()
print(json.dumps({"soup": n_tri, "gen_s": t1 - t0, "build_s": t2 - t1, "wide4_nodes": scene.info["n_wide_nodes"], "depth4": scene.info["wide_depth"]}), flush=True)
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev); g.manual_seed(0x5EED0002)
o = torch.rand((n_rays, 3), device=dev, generator=g)
d = torch.randn((n_rays, 3), device=dev, generator=g); d = d / d.norm(dim=1, keepdim=True)
q = torch.rand((n_rays, 3), device=dev, generator=g)
sets = {"closest": (torch.cat([o, d], 1).contiguous(), False), "shadow": (torch.cat([o, q - o], 1).contiguous(), True)}
hits = torch.empty((n_rays, 4), dtype=torch.float32, device=dev)
base_hits = {}
for v in variants:
    env = dict(kv.split("=") for kv in v.split(",") if kv)
    for k in [k for k in os.environ if k.startswith("FRT_") and k != "FRT_ASSETS"]:
        os.environ.pop(k, None)
    for k, val in env.items():
        os.environ["FRT_" + k.upper()] = val
    ctx = api.Context(devices=[0])
    t0 = time.time(); ctx.upload(scene); up = time.time() - t0
    for label, (rays, any_hit) in sets.items():
        for sort in (False, True):
            torch.cuda.synchronize()
            ms = min(ctx.trace_device(rays.data_ptr(), n_rays, hits.data_ptr(), any_hit, sort=sort) for _ in range(3))
            ids = hits[:, 3].view(torch.int32).clone()
            key = label
            if key not in base_hits:
                base_hits[key] = ids
            same = float(((ids >= 0) == (base_hits[key] >= 0)).float().mean()) if any_hit else float((ids == base_hits[key]).float().mean())
            ctx.set_flags(api.FRT_FLAG_COUNTERS); sub = min(n_rays, 1 << 21)
            c0 = ctx.counters(); ctx.trace_device(rays.data_ptr(), sub, hits.data_ptr(), any_hit, sort=sort); c1 = ctx.counters()
            ctx.set_flags(0)
            nodes = (c1["node_visits"] - c0["node_visits"]) / sub; tris = (c1["tri_tests"] - c0["tri_tests"]) / sub
            nb = (96.0 if env.get("wide8_stride", "6") != "8" else 128.0) if env.get("wide8", "1") != "0" and n_tri >= 65536 else 128.0
            b_ray = nb * nodes + 48 * tris + 40
            gbs = n_rays * b_ray / (ms * 1e-3) / 1e9
            print(json.dumps({"variant": v, "query": label, "sorted": sort, "Mrays_per_s": n_rays / ms / 1e3, "ms": ms, "same_as_first_variant": same,
                              "hit_fraction": float((ids >= 0).float().mean()), "nodes_per_ray": nodes, "tris_per_ray": tris, "bytes_per_ray": b_ray,
                              "algorithmic_GBs": gbs, "frac_of_hbm_peak": gbs / PEAK, "upload_s": up}), flush=True)
    ctx.close()
