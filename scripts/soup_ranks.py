"""GPU box with N GPUs: does the soup cast slow down when several GPUs cast at once?  One process per GPU (spawned
here, no torchrun): each builds the soup, then casts (a) alone, one rank after the other, and (b) all at the same time.
  python scripts/soup_ranks.py N [n_tri] [n_rays_per_gpu]"""
import os, sys, time, json, subprocess
HERE = os.path.dirname(os.path.abspath(__file__))
if len(sys.argv) > 1 and sys.argv[1] == "--child":
    rank, world, n_tri, n = int(sys.argv[2]), int(sys.argv[3]), int(float(sys.argv[4])), int(sys.argv[5])
    seed = int(os.environ.get("SOUP_SEED", rank))
    sys.path.insert(0, os.path.dirname(HERE))
    import torch
    from first_raytracer_b200 import api
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    scene = api.HostScene(soup=api.soup_triangles(n_tri), fast_bvh=True, max_leaf=4)
    ctx = api.Context(devices=[rank]); ctx.upload(scene)
    g = torch.Generator(device=dev); g.manual_seed(0x5EED0002 + seed)
    o = torch.rand((n, 3), device=dev, generator=g); d = torch.randn((n, 3), device=dev, generator=g); d = d / d.norm(dim=1, keepdim=True)
    rays = torch.cat([o, d], 1).contiguous(); hits = torch.empty((n, 4), dtype=torch.float32, device=dev)
    ctx.trace_device(rays.data_ptr(), n, hits.data_ptr())
    # plain HBM copy bandwidth of this device (2 GiB, read + write) and the PCIe rate of a pinned 256 MiB buffer
    a = torch.empty(1 << 31, dtype=torch.uint8, device=dev); b2 = torch.empty_like(a)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    b2.copy_(a); e0.record(); b2.copy_(a); e1.record(); e1.synchronize(); copy_gbs = 2 * (1 << 31) / e0.elapsed_time(e1) / 1e6
    h = torch.empty(1 << 28, dtype=torch.uint8).pin_memory()
    a[: 1 << 28].copy_(h, non_blocking=True); e0.record(); a[: 1 << 28].copy_(h, non_blocking=True); e1.record(); e1.synchronize()
    h2d_gbs = (1 << 28) / e0.elapsed_time(e1) / 1e6
    del a, b2
    flag = sys.argv[6]

    def wait(name):
        while not os.path.exists(f"{flag}.{name}"):
            time.sleep(0.01)
    open(f"{flag}.ready{rank}", "w").close()
    wait(f"solo{rank}")
    solo = [ctx.trace_device(rays.data_ptr(), n, hits.data_ptr()) for _ in range(3)]
    open(f"{flag}.solodone{rank}", "w").close()
    wait("all")
    together = [ctx.trace_device(rays.data_ptr(), n, hits.data_ptr()) for _ in range(6)]
    smi = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,power.draw,clocks_event_reasons.sw_power_cap", "--format=csv,noheader", "-i", str(rank)], capture_output=True, text=True).stdout.strip()
    print(json.dumps({"rank": rank, "solo_ms": [round(x, 2) for x in solo], "together_ms": [round(x, 2) for x in together],
                      "solo_Mrays": round(n / min(solo) / 1e3), "together_Mrays": round(n / min(together) / 1e3), "smi_after": smi, "seed": seed,
                      "hbm_copy_GBs": round(copy_gbs), "h2d_GBs": round(h2d_gbs, 1)}), flush=True)
    sys.exit(0)
N = int(sys.argv[1]); n_tri = sys.argv[2] if len(sys.argv) > 2 else "1e7"; n = sys.argv[3] if len(sys.argv) > 3 else str(1 << 24)
flag = f"/tmp/soup_ranks_{os.getpid()}"
procs = [subprocess.Popen([sys.executable, __file__, "--child", str(r), str(N), n_tri, n, flag], stdout=subprocess.PIPE, text=True) for r in range(N)]
def wait(name):
    while not os.path.exists(f"{flag}.{name}"):
        time.sleep(0.05)
for r in range(N): wait(f"ready{r}")
for r in range(N):
    open(f"{flag}.solo{r}", "w").close(); wait(f"solodone{r}")
open(f"{flag}.all", "w").close()
for p in procs:
    out, _ = p.communicate(timeout=600)
    print([l for l in out.splitlines() if l.startswith("{")][-1])
