import os, sys, time
sys.path.insert(0, "/root/repo")
os.environ["FRT_HOST_TIMING"] = "1"
from first_raytracer_b200 import api
import torch
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1
scene = api.HostScene("cornell", 16 / 9, 2)
for i in range(3):
    t = time.perf_counter(); img, st, secs = scene.render(1920, 1080, 64, n_gpus=n); dt = time.perf_counter() - t
    print(f"render {i}: python wall {dt*1e3:.1f} ms, Render() {secs*1e3:.1f} ms, device {st['render_ms']:.1f} ms", flush=True)
