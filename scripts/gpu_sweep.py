"""GPU box: sweep the CTA-resident wavefront's launch knobs on the bench workload at reduced spp.
usage: python scripts/gpu_sweep.py "cfg:slots:refill[:leaf[:sort]]" ...   (debug numbers, not bench values)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from first_raytracer_b200 import api

NX, NY = 1920, 1080
SPP = int(os.environ.get("SPP", "32"))
SCENE = os.environ.get("SCENE", "cornell")
scenes = {}
for spec in sys.argv[1:]:
    f = spec.split(":")
    cfg, slots, refill = f[0], f[1], f[2]
    leaf = int(f[3]) if len(f) > 3 else 2
    os.environ["FRT_WF_SORT"] = f[4] if len(f) > 4 else "1"
    os.environ["FRT_WF_CONFIG"], os.environ["FRT_WF_SLOTS"], os.environ["FRT_WF_REFILL"] = cfg, slots, refill
    if leaf not in scenes:
        scenes[leaf] = api.HostScene(SCENE, NX / NY, leaf)
    try:
        ctx = api.Context(devices=[0])
        ctx.upload(scenes[leaf])
        best = None
        for rep in range(3):
            ctx.clear_film()
            ctx.render_tile(NX, NY, 0, SPP, seed=1)
            _, st = ctx.readback(want_image=False)
            best = st if best is None or st["render_ms"] < best["render_ms"] else best
        st = best
        ms = st["render_ms"]
        print(f"cfg={cfg} slots={slots} refill={refill} leaf={leaf} sort={os.environ['FRT_WF_SORT']}: {NX*NY*SPP/ms/1e3:8.1f} Msamples/s  "
              f"{(st['closest_rays']+st['shadow_rays'])/ms/1e3:8.1f} Mrays/s  {ms:7.2f} ms  gen {st['generate_ms']:.1f} "
              f"ext {st['extend_ms']:.1f} shade {st['shade_ms']:.1f} con {st['connect_ms']:.1f} iters {st['iterations']}", flush=True)
        ctx.close()
    except Exception as exc:
        print(f"cfg={cfg} slots={slots} refill={refill}: FAILED {exc}", flush=True)
