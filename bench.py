"""bench.py — BASELINE.json's metric on BASELINE.json's config (configs[1]):
Cornell box 1920x1080, 1024 spp, NEE+MIS, samples/s and Mrays/s at N B200s.

  python bench.py --gpus N --steps K --warmup W            (own arm: CUDA wavefront integrator)
  python bench.py --impl reference --gpus N --steps K ...  (the reference's own CPU path::Li, oracle/_ref)

One step = one full render of the workload (all 1024 spp of every pixel).  With N ranks the spp
range is split evenly (strong scaling: total work fixed), each rank renders its share on its own
GPU and the partial radiance films are summed with ONE NCCL reduce; that reduce is inside the
timed step.  Timing: CUDA events on the stream the kernels are launched on, per step, max over
ranks of the summed step times; an L2 flush (256 MiB memset) runs between steps, untimed.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parent
sys.path.insert(0, str(REPO))

NX, NY, SPP = 1920, 1080, 1024
SCENE = "cornell"
WORKLOAD = f"Cornell box (CornellBox-Original.obj, 36 triangles) {NX}x{NY} {SPP} spp NEE+MIS depth<=34"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="frt", choices=["frt", "reference"])
    ap.add_argument("--spp", type=int, default=SPP, help="debug only: a run with --spp != 1024 is not a bench value")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-roofline", action="store_true", help="skip the per-kernel timing pass (profiling runs)")
    ap.add_argument("--scene", default="cornell", choices=["cornell", "veach", "teapot"],
                    help="other BASELINE.json configs at N GPUs (veach: config 3, teapot: config 4, 256 spp); "
                         "the contract bench is the default")
    ap.add_argument("--streamed-pool", action="store_true",
                    help="A/B only: the first implementation (global-memory path pool, 4 kernels per iteration)")
    return ap.parse_args()


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], threading.Event()

    def run(self):
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([c.strip() for c in out.strip().split(",")])
            except Exception:
                pass
            self.stop_flag.wait(0.2)

    def summary(self):
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(len(r) >= 6 and r[2 + i].lower() == "active" for r in self.rows)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


# ---------------------------------------------------------------------------------------------
# CPU side: the reference's own path::Li (oracle/_ref/frt_ref, compiled from /root/reference by
# oracle/Makefile).  Used only as the timed baseline, never on the product path.
# ---------------------------------------------------------------------------------------------
def cpu_reference_sample(spp: int = 1, ystep: int = 1, threads: int | None = None):
    """One bounded sample of the workload on the host cores: full 1920x1080 frame (every
    `ystep`-th row) at `spp` samples per pixel through the reference's taskflow row loop."""
    sys.path.insert(0, str(REPO / "tests"))
    import refutil
    if not refutil.REF_BIN.exists():
        raise RuntimeError("oracle/_ref/frt_ref is missing (build() compiles it where /root/reference exists)")
    cores = threads or os.cpu_count() or 1
    _, info = refutil.ref_render(SCENE, NX, NY, spp, ctr=False, seed=1, threads=cores, ystep=ystep)
    return info, cores


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    # size one step at ~5-10 s of CPU work: a quick probe at 1/16 of the rows first
    probe, cores = cpu_reference_sample(spp=1, ystep=16)
    rate = probe["samples_per_s"]
    target = 8.0 * rate  # samples in ~8 s
    ystep = max(1, int(round(NX * NY / target)))
    times, samples, rays = [], 0.0, 0.0
    for i in range(args.warmup + args.steps):
        info, cores = cpu_reference_sample(spp=1, ystep=ystep)
        if i >= args.warmup:
            times.append(info["seconds"])
            samples += info["samples"]
            rays += info["closest_rays"] + info["shadow_rays"]
    total = sum(times)
    value = samples / total / 1e6
    sample = f"{NX}x{NY} every {ystep}th row x 1 spp per step ({int(samples / len(times))} samples), taskflow {cores} threads"
    line = {
        "impl": "reference", "metric": "samples/s (1080p Cornell NEE+MIS)", "value": value, "unit": "Msamples/s",
        "mrays_per_s": rays / total / 1e6, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * total / len(times), "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": WORKLOAD, "note": "reference CPU path::Li timed on a bounded sample; samples/s is linear in spp"},
        "cpu_baseline": {"value": value, "unit": "Msamples/s", "cores": cores, "kind": "reference", "sample": sample},
        "e2e": {"value": value, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)
    return 0


# ---------------------------------------------------------------------------------------------
# Own arm
# ---------------------------------------------------------------------------------------------
class _DevBuf:
    """__cuda_array_interface__ view of a device pointer owned by libfrt (the film)."""
    def __init__(self, ptr: int, n: int):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f4", "data": (ptr, False), "version": 2}


def scene_bytes(scene) -> int:
    d = scene.desc
    n = d.n_nodes * 128 + d.n_tris * 48 + d.n_spheres * 32 + d.n_materials * 64 + d.n_lights * 128 + 100
    if d.tri_normals:
        n += d.n_tris * 48
    return int(n)


def algorithmic_bytes(kernel: str, st: dict, per_ray_nodes: float, per_ray_tris: float) -> float:
    """Algorithmic bytes moved by all launches of `kernel` in one render:
    extend : per closest ray 32 B ray in + 16 B hit out + 128 B per wide node + 48 B per triangle
    connect: per shadow ray 36 B segment+contribution in + 4 B queue entry + 16 B radiance RMW x2 + nodes/tris
    shade  : per path vertex 80 B state in (ray_o, ray_d, hit, thr, rad) + 64 B state out + 36 B shadow record
             + 8 B queue entries
    generate: per sample 4 B queue entry in + 80 B state out"""
    c, s, n = st["closest_rays"], st["shadow_rays"], st["samples"]
    geo = 128.0 * per_ray_nodes + 48.0 * per_ray_tris
    if kernel == "extend":
        return c * (48.0 + geo)
    if kernel == "connect":
        return s * (72.0 + geo)
    if kernel == "shade":
        return c * 144.0 + s * 40.0 + n * 4.0
    return n * 84.0


def run_frt(args):
    import torch
    import torch.distributed as dist
    from first_raytracer_b200 import api

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device (the product has no CPU path; use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    from first_raytracer_b200 import sharding
    spp = args.spp
    s0, s1 = sharding.spp_share(0, spp, rank, world)

    scene = api.HostScene(SCENE, NX / NY, int(os.environ.get("FRT_MAX_LEAF", "2")))
    ctx = api.Context(devices=[local])
    ctx.upload(scene)
    base_flags = api.FRT_FLAG_STREAMED_POOL if args.streamed_pool else 0
    ctx.set_flags(base_flags)
    stream = torch.cuda.ExternalStream(ctx.stream_handle(0), device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    film_t = None

    def one_step(timed: bool):
        nonlocal film_t
        flush.fill_(rank & 0xFF)  # L2 flush, untimed
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        ctx.clear_film()
        if s1 > s0:
            ctx.render_tile(NX, NY, s0, s1, seed=1)
        if world > 1:
            if film_t is None:
                p, n = ctx.film_device_ptr(0)
                film_t = torch.as_tensor(_DevBuf(p, n), device=dev)
            with torch.cuda.stream(stream):
                sharding.reduce_film(film_t, dst=0)  # the one collective of the path: sum of partial films
        e1.record(stream)
        e1.synchronize()
        _, st = ctx.readback(want_image=False)
        return e0.elapsed_time(e1), st

    for _ in range(args.warmup):
        one_step(False)
    sampler = ClockSampler(local)
    sampler.start()
    times, launches, closest, shadow, samples = [], 0, 0, 0, 0
    for _ in range(args.steps):
        ms, st = one_step(True)
        times.append(ms)
        launches += st["kernel_launches"] + (1 if world > 1 else 0)
        closest += st["closest_rays"]; shadow += st["shadow_rays"]; samples += st["samples"]
    sampler.stop_flag.set()
    sampler.join()
    total_ms = sum(times)
    # max over ranks of the device time; sums over ranks of the work
    if world > 1:
        t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms = float(t.item())
        w = torch.tensor([launches, closest, shadow, samples], dtype=torch.float64, device=dev)
        dist.all_reduce(w, op=dist.ReduceOp.SUM)
        launches, closest, shadow, samples = (int(x) for x in w.tolist())
    result = {"total_ms": total_ms, "launches": launches, "closest": closest, "shadow": shadow, "samples": samples,
              "clocks": sampler.summary(), "last_stats": st}
    return result, ctx, scene, (rank, world, local, dev), base_flags, stream


def roofline_pass(ctx, api, spp_range, peaks, base_flags, stream):
    """Dominant kernel = k_wavefront (the whole tile is one persistent launch).  Its launch duration
    is measured live with CUDA events on the launching stream over one more full render of this
    rank's share; node / triangle fetches per ray come from a short counting render
    (FRT_FLAG_COUNTERS, 16 spp).  achieved = rays x B_ray / duration with SURVEY 8(d)'s
    B_ray = 128 N_node + 48 N_tri + 32 + 16.  With --streamed-pool the top kernel of the four is
    reported instead (per-kernel events, FRT_FLAG_KERNEL_TIMERS)."""
    import torch
    s0, s1 = spp_range
    ctx.set_flags(base_flags | api.FRT_FLAG_COUNTERS)
    ctx.clear_film()
    ctx.render_tile(NX, NY, 0, 16, seed=1)
    _, sc = ctx.readback(want_image=False)
    rays = max(1, sc["closest_rays"] + sc["shadow_rays"])
    per_nodes, per_tris = sc["node_visits"] / rays, sc["tri_tests"] / rays
    streamed = bool(base_flags & api.FRT_FLAG_STREAMED_POOL)
    ctx.set_flags(base_flags | (api.FRT_FLAG_KERNEL_TIMERS if streamed else 0))
    ctx.clear_film()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    ctx.render_tile(NX, NY, s0, s1, seed=1)
    e1.record(stream)
    e1.synchronize()
    _, st = ctx.readback(want_image=False)
    ctx.set_flags(base_flags)
    kms = {k: st[k + "_ms"] for k in ("generate", "extend", "shade", "connect")}
    peak = peaks.get("hbm_gbs", 6650.0)
    b_ray = 128.0 * per_nodes + 48.0 * per_tris + 48.0
    if streamed:
        top = max(kms, key=kms.get)
        n_launch = max(1, st["iterations"])
        avg_ms = kms[top] / n_launch
        bytes_per_launch = algorithmic_bytes(top, st, per_nodes, per_tris) / n_launch
        kernel = "k_" + top
    else:
        n_launch, avg_ms, kernel = 1, e0.elapsed_time(e1), "k_wavefront"
        bytes_per_launch = (st["closest_rays"] + st["shadow_rays"]) * b_ray
    achieved = bytes_per_launch / (avg_ms * 1e-3) / 1e9 if avg_ms > 0 else 0.0
    # dram__bytes_read.sum + dram__bytes_write.sum of this very launch (k_wavefront, 1080p x 1024 spp on one GPU) from the
    # ncu pass of scripts/gpu_traffic.sh: profiles/r01_ncu_traffic_1024spp.csv.  Other launch shapes: not captured.
    traffic = 27660544 + 294398976 if (not streamed and (s1 - s0) == 1024) else None
    return {
        "bound": "hbm", "kernel": kernel, "achieved": achieved, "peak": peak, "unit": "GB/s",
        "frac": achieved / peak, "traffic": traffic,
        "peak_source": "MEASURED_PEAKS.json (measured)" if "hbm_gbs" in peaks else "fallback 6650 GB/s",
        "avg_launch_ms": avg_ms, "launches": n_launch, "algorithmic_bytes_per_launch": bytes_per_launch,
        "stage_ms_mean_over_ctas" if not streamed else "kernel_ms": kms,
        "stage_share": {k: v / max(1e-9, sum(kms.values())) for k, v in kms.items()},
        "nodes_per_ray": per_nodes, "tris_per_ray": per_tris, "bytes_per_ray": b_ray,
        "wavefront_iterations": st["iterations"],
        # one ncu --set full capture of a 1080p x 8 spp launch (profiles/r01_ncu_wavefront.txt): DRAM bytes of that
        # launch next to its algorithmic bytes; the scene and the path pool never leave L1/L2/shared memory
        "traffic_ncu_8spp_launch": {"dram_bytes": 25.35e6 + 5.72e6, "algorithmic_bytes": 111.2e6 * b_ray},
        "note": "scene (36 triangles) is L1/L2 resident: the HBM fraction is reported as the contract asks; "
                "the binding resource is SM issue slots (profiles/)",
    }


def e2e_pass(scene, steps, n_gpus, spp):
    """The call a user of the reference makes: renderer<gpu_path_integrator>::Render(scene, film)
    (frth_render) with HOST scene in and HOST film out — context creation, scene flatten + upload,
    render, multi-GPU film sum, device->host film copy and viewer::add_sample all inside."""
    scene.render(64, 64, 4, n_gpus=n_gpus)  # CUDA context / module load warm-up
    t0 = time.perf_counter()
    for _ in range(steps):
        img, st, secs = scene.render(NX, NY, spp, seed=1, n_gpus=n_gpus)
    dt = time.perf_counter() - t0
    return steps * NX * NY * spp / dt / 1e6, float(img.mean()), dt / steps


_JSON_OUT = None


def emit(line: dict):
    _JSON_OUT.write(json.dumps(line) + "\n")
    _JSON_OUT.flush()


def main():
    global SCENE, SPP, WORKLOAD, _JSON_OUT
    args = parse()
    # stdout carries exactly ONE line, the JSON: everything else written to fd 1 (renderer<>::Render prints the
    # reference's timing block with std::cout, integrator.h:33-43) goes to stderr
    sys.stdout.flush()
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    if args.scene != "cornell":
        SCENE, SPP = args.scene, 256
        WORKLOAD = {"veach": f"veach_mi (GGX rough_conductor plates, 5 sphere lights) {NX}x{NY} {SPP} spp NEE+MIS",
                    "teapot": f"mesh scene (teapot stand-in, 15 704 triangles) {NX}x{NY} {SPP} spp NEE+MIS"}[SCENE]
        if args.spp == 1024:
            args.spp = SPP
    if args.impl == "reference":
        return run_reference(args)
    from first_raytracer_b200 import api
    res, ctx, scene, (rank, world, local, dev), base_flags, stream = run_frt(args)
    import torch
    import torch.distributed as dist
    peaks = {}
    try:
        peaks = json.loads((REPO / "MEASURED_PEAKS.json").read_text())
    except Exception:
        pass
    spp = args.spp
    steps = args.steps
    total_s = res["total_ms"] * 1e-3
    value = steps * NX * NY * spp / total_s / 1e6
    roof = cpu = e2e = None
    if rank == 0 and not args.no_roofline:
        roof = roofline_pass(ctx, api, (0, max(1, spp // world)), peaks, base_flags, stream)
    ctx.close()
    # The end-to-end leg runs renderer<gpu_path_integrator> with n_gpus = world inside rank 0's process; the other
    # ranks must leave their GPUs idle meanwhile, so they wait on a CPU (gloo) barrier: an NCCL barrier spins a
    # kernel on the waiting GPU.
    cpu_group = dist.new_group(backend="gloo") if world > 1 else None
    if world > 1:
        torch.cuda.synchronize()
        dist.barrier(group=cpu_group)
    if rank == 0 and not args.no_e2e:
        e2e_val, e2e_mean, e2e_s = e2e_pass(scene, max(1, min(steps, 3)), world, spp)
        e2e = {"value": e2e_val, "unit": "Msamples/s", "h2d_bytes_per_step": scene_bytes(scene),
               "d2h_bytes_per_step": NX * NY * 3 * 4 + 128 * world, "seconds_per_step": e2e_s,
               "mean_radiance": e2e_mean,
               "call": "renderer<gpu_path_integrator>::Render via frth_render (host scene in, host film out)"}
    if world > 1:
        dist.barrier(group=cpu_group)
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            probe, cores = cpu_reference_sample(spp=1, ystep=16)
            ystep = max(1, int(round(NX * NY / (15.0 * probe["samples_per_s"]))))
            info, cores = cpu_reference_sample(spp=1, ystep=ystep)
            cpu = {"value": info["samples_per_s"] / 1e6, "unit": "Msamples/s", "cores": cores, "kind": "reference",
                   "mrays_per_s": info["mrays_per_s"], "seconds": info["seconds"],
                   "sample": f"{NX}x{NY} every {ystep}th row x 1 spp ({int(info['samples'])} samples) through the "
                             f"reference's path::Li on its taskflow, {cores} threads"}
        except Exception as exc:  # the baseline is reported, never required
            cpu = {"value": None, "unit": "Msamples/s", "cores": 0, "kind": "reference", "sample": f"failed: {exc}"}
    if rank == 0:
        line = {
            "metric": "samples/s (1080p Cornell NEE+MIS)" if SCENE == "cornell" else f"samples/s (1080p {SCENE} NEE+MIS)",
            "value": value, "unit": "Msamples/s",
            "mrays_per_s": (res["closest"] + res["shadow"]) / total_s / 1e6,
            "rays_per_sample": (res["closest"] + res["shadow"]) / max(1, res["samples"]),
            "n_gpus": world, "steps": steps, "warmup": args.warmup, "ms_per_step": res["total_ms"] / steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": WORKLOAD if spp == SPP else WORKLOAD.replace(f"{SPP} spp", f"{spp} spp (DEBUG, not a bench value)"),
                       "parallelism": f"spp range split over {world} rank(s), one NCCL reduce of the film",
                       "l2": "256 MiB memset between steps (untimed)",
                       "pool": "streamed (global memory, 4 kernels/iteration)" if args.streamed_pool else "CTA-resident (shared memory, one persistent kernel per tile)",
                       "timing": "CUDA events on the launching stream per step, summed, max over ranks"},
            "clocks": res["clocks"], "e2e": e2e, "gpu_launches": res["launches"],
            "roofline": roof, "cpu_baseline": cpu,
        }
        emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
