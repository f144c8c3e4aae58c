"""GPU box: the BASELINE.json configs other than the headline one, on ONE GPU (debug harness, prints one
JSON line per config; the contract bench is bench.py).
  config 3  veach_mi (rough_conductor plates, 5 sphere lights) 768x512 and 1920x1080, 256 spp
  config 4  mesh scene (teapot stand-in, 15 704 triangles) 1920x1080, 256 spp
  config 5  synthetic triangle soup: incoherent closest-hit and shadow ray casts, unsorted / Morton-sorted
usage: python scripts/bench_scenes.py [veach] [teapot] [hdr] [soup]      env: SOUP_TRIS (1e7), SOUP_RAYS (2**26)"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from first_raytracer_b200 import api

PEAK = 6546.2
try:
    PEAK = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
which = sys.argv[1:] or ["veach", "teapot", "soup"]


def render_cfg(name, nx, ny, spp, leaf=2):
    scene = api.HostScene(name, nx / ny, leaf)
    ctx = api.Context(devices=[0])
    ctx.upload(scene)
    best = None
    for _ in range(3):
        ctx.clear_film()
        ctx.render_tile(nx, ny, 0, spp, seed=1)
        img, st = ctx.readback()
        best = st if best is None or st["render_ms"] < best["render_ms"] else best
    ctx.set_flags(api.FRT_FLAG_COUNTERS)
    ctx.clear_film(); ctx.render_tile(nx, ny, 0, 4, seed=1); _, sc = ctx.readback(want_image=False)
    ctx.close()
    rays = best["closest_rays"] + best["shadow_rays"]; crays = max(1, sc["closest_rays"] + sc["shadow_rays"])
    ms = best["render_ms"]
    print(json.dumps({"config": f"{name} {nx}x{ny} {spp} spp", "Msamples_per_s": nx * ny * spp / ms / 1e3,
                      "Mrays_per_s": rays / ms / 1e3, "ms": ms, "rays_per_sample": rays / (nx * ny * spp),
                      "nodes_per_ray": sc["node_visits"] / crays, "tris_per_ray": sc["tri_tests"] / crays,
                      "mean_radiance": float(img.mean() / spp), "nonfinite_free": bool(np.isfinite(img).all()),
                      "stage_ms": {k: best[k + "_ms"] for k in ("generate", "extend", "shade", "connect")}}), flush=True)


def morton_order(o):
    q = (o.clamp(0, 1 - 1e-7) * 1024).to(torch.int64)
    def spread(v):
        v = (v | (v << 16)) & 0x030000FF; v = (v | (v << 8)) & 0x0300F00F
        v = (v | (v << 4)) & 0x030C30C3; v = (v | (v << 2)) & 0x09249249
        return v
    return torch.argsort(spread(q[:, 0]) | (spread(q[:, 1]) << 1) | (spread(q[:, 2]) << 2))


def soup():
    n_tri = int(float(os.environ.get("SOUP_TRIS", "1e7"))); n_rays = int(os.environ.get("SOUP_RAYS", str(2 ** 26)))
    t0 = time.time(); tri = api.soup_triangles(n_tri); t1 = time.time()
    scene = api.HostScene(soup=tri, fast_bvh=True, max_leaf=int(os.environ.get("SOUP_LEAF", "2"))); t2 = time.time()
    info = scene.info() if callable(scene.info) else scene.info
    ctx = api.Context(devices=[0]); ctx.upload(scene); t3 = time.time()
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev); g.manual_seed(0x5EED0002)
    o = torch.rand((n_rays, 3), device=dev, generator=g)
    d = torch.randn((n_rays, 3), device=dev, generator=g); d = d / d.norm(dim=1, keepdim=True)
    q = torch.rand((n_rays, 3), device=dev, generator=g)
    sets = {"closest": (torch.cat([o, d], 1).contiguous(), False), "shadow": (torch.cat([o, q - o], 1).contiguous(), True)}
    order = morton_order(o)
    hits = torch.empty((n_rays, 4), dtype=torch.float32, device=dev)
    base = {"config": f"soup {n_tri} triangles, {n_rays} rays", "gen_s": t1 - t0, "bvh_build_s": t2 - t1, "upload_s": t3 - t2,
            "wide_nodes": info["n_wide_nodes"], "depth": info["wide_depth"]}
    for label, (rays, any_hit) in sets.items():
        for variant, r in (("unsorted", rays), ("morton", rays[order].contiguous())):
            torch.cuda.synchronize()
            ms = min(ctx.trace_device(r.data_ptr(), n_rays, hits.data_ptr(), any_hit) for _ in range(3))
            hit_frac = float((hits[:, 3].view(torch.int32) >= 0).float().mean())
            # node / triangle fetches per ray from a counting build of the same kernel on a 4M-ray prefix
            ctx.set_flags(api.FRT_FLAG_COUNTERS); sub = min(n_rays, 1 << 22)
            c0 = ctx.counters()
            ctx.trace_device(r.data_ptr(), sub, hits.data_ptr(), any_hit)
            c1 = ctx.counters()
            ctx.set_flags(0)
            nodes = (c1["node_visits"] - c0["node_visits"]) / sub; tris = (c1["tri_tests"] - c0["tri_tests"]) / sub
            b_ray = 128 * nodes + 48 * tris + 24 + 16  # node + triangle fetches + ray in (24 B here) + hit out
            gbs = n_rays * b_ray / (ms * 1e-3) / 1e9
            print(json.dumps({**base, "query": label, "order": variant, "Mrays_per_s": n_rays / ms / 1e3, "ms": ms,
                              "hit_fraction": hit_frac, "nodes_per_ray": nodes, "tris_per_ray": tris, "bytes_per_ray": b_ray,
                              "roofline": {"bound": "hbm", "achieved": gbs, "peak": PEAK, "unit": "GB/s", "frac": gbs / PEAK}}),
                  flush=True)
    ctx.close()


for w in which:
    if w == "veach":
        render_cfg("veach", 768, 512, 256); render_cfg("veach", 1920, 1080, 256)
    elif w == "teapot":
        render_cfg("teapot", 1920, 1080, 256)
    elif w == "hdr":  # the same scenes lit by the HDR lat-long environment image
        render_cfg("teapot+hdr", 1920, 1080, 256); render_cfg("veach+hdr", 1920, 1080, 256)
    elif w == "cornell":
        render_cfg("cornell", 512, 512, 256)
    elif w == "soup":
        soup()
