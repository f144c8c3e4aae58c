"""GPU box: wavefront kernel variants on the render configs, one JSON line each.
  python scripts/wf_sweep.py "scene:spp" ... -- "ENV=VAL,ENV=VAL" "..."     (env names without the FRT_ prefix)"""
import json, os, subprocess, sys
HERE = os.path.dirname(os.path.abspath(__file__))
if len(sys.argv) > 1 and sys.argv[1] == "--child":
    sys.path.insert(0, os.path.dirname(HERE))
    from first_raytracer_b200 import api
    name, spp = sys.argv[2].split(":"); spp = int(spp)
    nx, ny = 1920, 1080
    scene = api.HostScene(name, nx / ny, int(os.environ.get('SWEEP_LEAF', '2')))
    ctx = api.Context(devices=[0]); ctx.upload(scene)
    best = None
    for _ in range(3):
        ctx.clear_film(); ctx.render_tile(nx, ny, 0, spp, seed=1)
        img, st = ctx.readback()
        best = st if best is None or st["render_ms"] < best["render_ms"] else best
    ms = best["render_ms"]
    print(json.dumps({"scene": name, "spp": spp, "env": sys.argv[3], "Msamples_per_s": round(nx * ny * spp / ms / 1e3, 1), "ms": round(ms, 2),
                      "mean": float(img.mean() / spp), "stage_ms": {k: round(best[k + "_ms"], 1) for k in ("generate", "extend", "shade", "connect")},
                      "iters": best["iterations"]}), flush=True)
    ctx.close()
    sys.exit(0)
args = sys.argv[1:]
cut = args.index("--")
scenes, variants = args[:cut], args[cut + 1:]
for v in variants:
    env = {k: val for k, val in os.environ.items() if not k.startswith("FRT_") or k == "FRT_ASSETS"}
    for kv in v.split(","):
        if kv:
            k, val = kv.split("="); env["FRT_" + k.upper()] = val
    for sc in scenes:
        r = subprocess.run([sys.executable, __file__, "--child", sc, v], env=env, capture_output=True, text=True, timeout=600)
        out = [l for l in r.stdout.splitlines() if l.startswith("{")]
        print(out[-1] if out else json.dumps({"scene": sc, "env": v, "failed": r.stderr[-300:]}), flush=True)
