"""bench.py — BASELINE.json's metric on BASELINE.json's config (configs[1]):
Cornell box 1920x1080, 1024 spp, NEE+MIS, samples/s and Mrays/s at N B200s.

  python bench.py --gpus N --steps K --warmup W            (own arm: CUDA wavefront integrator)
  python bench.py --impl reference --gpus N --steps K ...  (the reference's own CPU path::Li, oracle/_ref)

One step = one full render of the workload (all 1024 spp of every pixel).  With N ranks the spp
range is split evenly (strong scaling: total work fixed), each rank renders its share on its own
GPU and the partial radiance films are summed with ONE NCCL reduce; that reduce is inside the
timed step.  Timing: CUDA events on the stream the kernels are launched on, per step, max over
ranks of the summed step times; an L2 flush (256 MiB memset) runs between steps, untimed.

After the headline the same process measures BASELINE.json's other configs (the `configs` block of
the JSON line: config 1 Cornell 512^2 x 16 spp, config 3 veach_mi, config 4 mesh scene, config 5 the
10^7-triangle soup ray casts), each with its own clock summary and a parity bit; under torchrun every
rank takes its share of each (spp range / ray block), so SCALE carries them at N = 2, 4, 8 too.
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
GOLDEN = REPO / "tests" / "golden"


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
    ap.add_argument("--no-configs", action="store_true", help="skip BASELINE.json's other configs (profiling runs)")
    ap.add_argument("--soup-tris", type=float, default=1e7, help="debug only: config 5 is 10^7 triangles")
    ap.add_argument("--soup-rays", type=int, default=1 << 26, help="debug only: config 5 casts 2^26 rays")
    ap.add_argument("--scene", default="cornell", choices=["cornell", "veach", "teapot"],
                    help="headline scene (debug): the contract bench is the default")
    return ap.parse_args()


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int, period: float = 0.2):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag, self.period = index, [], threading.Event(), period

    def run(self):
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([c.strip() for c in out.strip().split(",")])
            except Exception:
                pass
            self.stop_flag.wait(self.period)

    def finish(self):
        self.stop_flag.set()
        self.join()
        return self.summary()

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
def cpu_reference_sample(spp: int = 1, ystep: int = 1, threads: int | None = None, scene: str | None = None,
                         nx: int = NX, ny: int = NY):
    """One bounded sample of the workload on the host cores: the full frame (every `ystep`-th row) at
    `spp` samples per pixel through the reference's taskflow row loop."""
    sys.path.insert(0, str(REPO / "tests"))
    import refutil
    if not refutil.REF_BIN.exists():
        raise RuntimeError("oracle/_ref/frt_ref is missing (build() compiles it where /root/reference exists)")
    cores = threads or os.cpu_count() or 1
    _, info = refutil.ref_render(scene or SCENE, nx, ny, spp, ctr=False, seed=1, threads=cores, ystep=ystep)
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


class Rig:
    """What every measurement below shares: rank layout, the frt context on this rank's GPU, the stream its kernels
    run on, the L2 flush buffer."""

    def __init__(self):
        import torch
        import torch.distributed as dist
        from first_raytracer_b200 import api, sharding
        self.torch, self.dist, self.api, self.sharding = torch, dist, api, sharding
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device (the product has no CPU path; use --impl reference for the CPU arm)")
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=self.dev)
        self.ctx = api.Context(devices=[self.local])
        self.stream = torch.cuda.ExternalStream(self.ctx.stream_handle(0), device=self.dev)
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=self.dev)
        self.film_t, self.film_key = None, None

    def barrier(self):
        self.torch.cuda.synchronize()
        if self.world > 1:
            self.dist.barrier()

    def reduce_film(self, nx, ny):
        """The one collective of the path: sum of the partial films onto rank 0 (NCCL over NVLink), on the render stream."""
        if self.world == 1:
            return
        if self.film_key != (nx, ny):
            p, n = self.ctx.film_device_ptr(0)
            self.film_t, self.film_key = self.torch.as_tensor(_DevBuf(p, n), device=self.dev), (nx, ny)
        with self.torch.cuda.stream(self.stream):
            self.sharding.reduce_film(self.film_t, dst=0)

    def render_step(self, nx, ny, s0, s1, seed=1):
        """One timed step: clear, render this rank's spp share, reduce.  → (ms on the launching stream, stats)."""
        torch = self.torch
        self.flush.fill_(self.rank & 0xFF)  # L2 flush, untimed
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(self.stream)
        self.ctx.clear_film()
        if s1 > s0:
            self.ctx.render_tile(nx, ny, s0, s1, seed=seed)
        elif self.world > 1:
            self.ctx.render_tile(nx, ny, 0, 0, seed=seed)  # allocates the film this rank contributes (zeros) to the reduce
        self.reduce_film(nx, ny)
        e1.record(self.stream)
        e1.synchronize()
        _, st = self.ctx.readback(want_image=False)
        return e0.elapsed_time(e1), st

    def per_rank(self, x: float) -> list:
        """x of every rank, in rank order (diagnostics: which rank sets the max)."""
        if self.world == 1:
            return [x]
        t = self.torch.zeros(self.world, dtype=self.torch.float64, device=self.dev)
        t[self.rank] = x
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return [float(v) for v in t.tolist()]

    def max_over_ranks(self, x: float) -> float:
        if self.world == 1:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(self, xs):
        if self.world == 1:
            return list(xs)
        t = self.torch.tensor(list(xs), dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return t.tolist()


def run_headline(rig: Rig, args, spp):
    s0, s1 = rig.sharding.spp_share(0, spp, rig.rank, rig.world)
    scene = rig.api.HostScene(SCENE, NX / NY, int(os.environ.get("FRT_MAX_LEAF", "2")))
    rig.ctx.upload(scene)
    rig.ctx.set_flags(0)
    for _ in range(args.warmup):
        rig.render_step(NX, NY, s0, s1)
    sampler = ClockSampler(rig.local)
    sampler.start()
    times, launches, closest, shadow, samples, st = [], 0, 0, 0, 0, None
    for _ in range(args.steps):
        ms, st = rig.render_step(NX, NY, s0, s1)
        times.append(ms)
        launches += st["kernel_launches"] + (1 if rig.world > 1 else 0)
        closest += st["closest_rays"]; shadow += st["shadow_rays"]; samples += st["samples"]
    clocks = sampler.finish()
    total_ms = rig.max_over_ranks(sum(times))
    launches, closest, shadow, samples = (int(x) for x in rig.sum_over_ranks([launches, closest, shadow, samples]))
    return {"total_ms": total_ms, "launches": launches, "closest": closest, "shadow": shadow, "samples": samples,
            "clocks": clocks, "last_stats": st}, scene, (s0, s1)


def ncu_metrics():
    """Counters one ncu capture of the dominant kernel gave (scripts/ncu_metrics.py writes the file from the .ncu-rep of
    scripts/gpu_profile.sh): DRAM bytes, L2 bytes, issue-slot utilisation, active lanes.  They belong to the capture's
    launch, not to this run, and are emitted with their source; absent file -> nulls."""
    p = REPO / "profiles" / "r02_ncu_metrics.json"
    try:
        return json.loads(p.read_text()), str(p.relative_to(REPO))
    except Exception:
        return None, None


def roofline_pass(rig: Rig, spp_range, peaks, clocks):
    """Dominant kernel = k_wavefront (the whole tile is one persistent launch).  Its launch duration is measured live
    with CUDA events on the launching stream over one more full render of this rank's share; node / triangle fetches
    per ray come from a short counting render (FRT_FLAG_COUNTERS, 16 spp).  achieved = rays x B_ray / duration with
    SURVEY 8(d)'s B_ray = 128 N_node + 48 N_tri + 32 + 16 — an ALGORITHMIC byte rate: the Cornell scene is 3.5 KB and
    lives in L1, so `frac` says how the fetch rate compares with HBM, not how much headroom there is; the binding
    resource is the SM's issue slots (`secondary`)."""
    torch, api, ctx = rig.torch, rig.api, rig.ctx
    s0, s1 = spp_range
    ctx.set_flags(api.FRT_FLAG_COUNTERS)
    ctx.clear_film()
    ctx.render_tile(NX, NY, 0, 16, seed=1)
    _, sc = ctx.readback(want_image=False)
    rays = max(1, sc["closest_rays"] + sc["shadow_rays"])
    per_nodes, per_tris = sc["node_visits"] / rays, sc["tri_tests"] / rays
    ctx.set_flags(0)
    ctx.clear_film()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(rig.stream)
    ctx.render_tile(NX, NY, s0, s1, seed=1)
    e1.record(rig.stream)
    e1.synchronize()
    _, st = ctx.readback(want_image=False)
    kms = {k: st[k + "_ms"] for k in ("generate", "extend", "shade", "connect")}
    peak = peaks.get("hbm_gbs", 6650.0)
    b_ray = 128.0 * per_nodes + 48.0 * per_tris + 48.0
    avg_ms = e0.elapsed_time(e1)
    n_rays = st["closest_rays"] + st["shadow_rays"]
    bytes_per_launch = n_rays * b_ray
    achieved = bytes_per_launch / (avg_ms * 1e-3) / 1e9 if avg_ms > 0 else 0.0
    m, src = ncu_metrics()
    traffic = secondary = l2 = None
    if m:
        k = m.get("k_wavefront", {})
        full = k.get("traffic")
        if full and full.get("dram_bytes") is not None and (s1 - s0) == SPP:
            # dram__bytes_read.sum + dram__bytes_write.sum of the very launch timed here (1080p x 1024 spp on one GPU),
            # from the ncu pass of scripts/gpu_traffic.sh; other launch shapes: not captured -> null
            traffic = full["dram_bytes"]
        if k.get("issue_active_pct") is not None:
            secondary = {"bound": "issue", "issue_active_frac": k["issue_active_pct"] / 100.0,
                         "active_lanes_per_inst": k.get("lanes_per_inst"), "lane_issue_frac": k["issue_active_pct"] / 100.0 * (k.get("lanes_per_inst") or 32) / 32.0,
                         "warp_inst_per_s": k.get("inst_executed", 0) / max(1e-9, k.get("duration_s", 1.0)),
                         "peak_warp_inst_per_s": 148 * 4 * (k.get("sm_mhz") or (clocks.get("sm_mhz") or 1965.0)) * 1e6,
                         "source": src, "capture": k.get("capture")}
        if k.get("lts_bytes") is not None:
            l2 = {"achieved_gbs": k["lts_bytes"] / max(1e-9, k.get("duration_s", 1.0)) / 1e9, "source": src}
    return {
        "bound": "hbm", "kernel": "k_wavefront", "achieved": achieved, "peak": peak, "unit": "GB/s",
        "frac": achieved / peak, "traffic": traffic, "traffic_source": src if traffic is not None else None,
        "peak_source": "MEASURED_PEAKS.json (measured)" if "hbm_gbs" in peaks else "fallback 6650 GB/s",
        "avg_launch_ms": avg_ms, "launches": 1, "algorithmic_bytes_per_launch": bytes_per_launch,
        "stage_ms_mean_over_ctas": kms,
        "stage_share": {k: v / max(1e-9, sum(kms.values())) for k, v in kms.items()},
        "nodes_per_ray": per_nodes, "tris_per_ray": per_tris, "bytes_per_ray": b_ray,
        "wavefront_iterations": st["iterations"], "secondary": secondary, "l2": l2,
        "note": "frac is an ALGORITHMIC byte rate over the HBM peak: the scene (36 triangles, 3.5 KB) is L1 resident and the "
                "real DRAM bytes are `traffic` (the film leaving L2); the bound that matters is `secondary` (SM issue slots)",
    }


# ---------------------------------------------------------------------------------------------
# BASELINE.json's other configs, measured in the same process (the `configs` block)
# ---------------------------------------------------------------------------------------------
def config_render(rig: Rig, name, nx, ny, spp, steps=3, fixture=None, fixture_block=1, leaf=0):
    """A render config on all ranks (spp range split, one NCCL reduce inside the step): Msamples/s from the max over
    ranks of the summed device step times; parity bit = mean radiance within 0.5 % of the as-shipped mt19937
    reference's (tests/golden/stat_*.npz, scripts/make_golden_stat.py) — rank 0 holds the reduced film."""
    api = rig.api
    scene = api.HostScene(name, nx / ny, leaf)
    rig.ctx.upload(scene)
    s0, s1 = rig.sharding.spp_share(0, spp, rig.rank, rig.world)
    rig.render_step(nx, ny, s0, s1)
    sampler = ClockSampler(rig.local, 0.1)
    sampler.start()
    ms_sum, rays, st = 0.0, 0, None
    for _ in range(steps):
        ms, st = rig.render_step(nx, ny, s0, s1)
        ms_sum += ms
        rays += st["closest_rays"] + st["shadow_rays"]
    clocks = sampler.finish()
    total_ms = rig.max_over_ranks(ms_sum)
    rays = int(rig.sum_over_ranks([rays])[0])
    out = {"workload": f"{name} {nx}x{ny} {spp} spp", "value": steps * nx * ny * spp / total_ms / 1e3, "unit": "Msamples/s",
           "mrays_per_s": rays / total_ms / 1e3, "ms_per_step": total_ms / steps, "steps": steps, "n_gpus": rig.world, "clocks": clocks}
    if fixture and rig.rank == 0:
        # the film on rank 0 after the last step's reduce
        sums, _ = rig.ctx.readback()
        mean = float(sums.mean() / spp)
        g = np.load(GOLDEN / f"{fixture}.npz")
        ref_mean = float(g["mean"])
        out["mean_radiance"] = mean
        out["parity"] = {"ok": bool(abs(mean - ref_mean) <= 5e-3 * ref_mean), "mean_rel_diff": abs(mean - ref_mean) / ref_mean,
                         "against": f"tests/golden/{fixture}.npz (reference path::Li, mt19937, mean {ref_mean:.6f})"}
    scene.close()
    return out


def config1(rig: Rig, want_cpu: bool):
    """Config 1 — Cornell-Original 512x512 16 spp, the one config the reference runs verbatim (main.cpp:222-252): the
    GPU value next to the reference's own seconds (timed live on this box's host cores when want_cpu, else the
    seconds recorded with the fixture in the build container)."""
    out = config_render(rig, "cornell", 512, 512, 16, steps=5, fixture="stat_cornell_512x512x1024")
    out["workload"] = "config 1: Cornell 512x512 16 spp"
    if rig.rank == 0:
        g = np.load(GOLDEN / "stat_cornell_512x512x16.npz")
        out["reference_cpu"] = {"seconds": float(g["seconds"]), "threads": int(g["threads"]), "seconds_1thread": float(g["seconds_1thread"]),
                                "where": "build container, scripts/make_golden_stat.py"}
        if want_cpu:
            try:
                info, cores = cpu_reference_sample(spp=16, scene="cornell", nx=512, ny=512)
                out["reference_cpu_live"] = {"seconds": info["seconds"], "threads": cores, "msamples_per_s": info["samples_per_s"] / 1e6,
                                             "mrays_per_s": info["mrays_per_s"], "where": "this box, oracle/_ref/frt_ref (path::Li on taskflow)"}
            except Exception as exc:
                out["reference_cpu_live"] = {"failed": str(exc)}
    return out


def config_soup(rig: Rig, n_tri: int, n_rays: int, peaks):
    """Config 5 — incoherent closest-hit + shadow casts on the synthetic soup, the ray array block-partitioned over the
    ranks (no exchange).  Device time = max over ranks of each rank's kernel (+ sort) time with rays resident in HBM;
    e2e = frt_trace_rays on this rank's block from PINNED host buffers (H2D, kernel, D2H inside, wall clock, max over
    ranks).  Parity bit: the Morton-sorted cast returns exactly the hits of the unsorted one (oracle parity of the same
    kernel at this size is tests/test_gpu_trace.py::test_ten_million_triangle_soup_parity)."""
    torch, api = rig.torch, rig.api
    t0 = time.time()
    tri = api.soup_triangles(n_tri)
    scene = api.HostScene(soup=tri, fast_bvh=True, max_leaf=int(os.environ.get("SOUP_LEAF", "3")))
    build_s = time.time() - t0
    del tri
    t0 = time.time()
    rig.ctx.upload(scene)
    upload_s = time.time() - t0
    b, e = rig.sharding.ray_share(n_rays, rig.rank, rig.world)
    n = e - b
    g = torch.Generator(device=rig.dev)
    g.manual_seed(0x5EED0002 + rig.rank)
    o = torch.rand((n, 3), device=rig.dev, generator=g)
    d = torch.randn((n, 3), device=rig.dev, generator=g)
    d = d / d.norm(dim=1, keepdim=True)
    q = torch.rand((n, 3), device=rig.dev, generator=g)
    sets = {"closest": (torch.cat([o, d], 1).contiguous(), False), "shadow": (torch.cat([o, q - o], 1).contiguous(), True)}
    del o, d, q
    hits = torch.empty((n, 4), dtype=torch.float32, device=rig.dev)
    hits2 = torch.empty((n, 4), dtype=torch.float32, device=rig.dev)
    peak = peaks.get("hbm_gbs", 6650.0)
    info = scene.info
    res = {"workload": f"config 5: soup {n_tri} triangles, {n_rays} rays block-partitioned over {rig.world} rank(s)", "n_gpus": rig.world,
           "bvh_build_s": build_s, "upload_s": upload_s, "wide4_nodes": info["n_wide_nodes"], "wide4_depth": info["wide_depth"], "casts": []}
    sampler = ClockSampler(rig.local, 0.1)
    sampler.start()
    for label, (rays, any_hit) in sets.items():
        for sort in (False, True):
            rig.barrier()
            ms = min(rig.ctx.trace_device(rays.data_ptr(), n, hits2.data_ptr() if sort else hits.data_ptr(), any_hit, sort=sort) for _ in range(3))
            ms_ranks = rig.per_rank(ms)
            ms = rig.max_over_ranks(ms)
            # node / triangle fetches per ray from a counting build of the same kernel on a prefix
            rig.ctx.set_flags(api.FRT_FLAG_COUNTERS)
            sub = min(n, 1 << 21)
            c0 = rig.ctx.counters()
            rig.ctx.trace_device(rays.data_ptr(), sub, hits2.data_ptr() if sort else hits.data_ptr(), any_hit, sort=sort)
            c1 = rig.ctx.counters()
            rig.ctx.set_flags(0)
            if sort:  # restore the full result the parity bit compares
                rig.ctx.trace_device(rays.data_ptr(), n, hits2.data_ptr(), any_hit, sort=True)
            nodes = (c1["node_visits"] - c0["node_visits"]) / sub
            tris = (c1["tri_tests"] - c0["tri_tests"]) / sub
            node_bytes = 96.0 if os.environ.get("FRT_WIDE8") == "1" else 128.0
            b_ray = node_bytes * nodes + 48.0 * tris + 24.0 + 16.0  # node + triangle fetches, ray in (24 B), hit out (16 B)
            gbs = n_rays * b_ray / (ms * 1e-3) / 1e9 / rig.world  # per GPU
            cast = {"query": label, "order": "morton (device sort inside)" if sort else "as given (incoherent)",
                    "value": n_rays / ms / 1e3, "unit": "Mrays/s", "ms": ms, "ms_per_rank": ms_ranks, "nodes_per_ray": nodes, "tris_per_ray": tris,
                    "bytes_per_ray": b_ray, "roofline": {"bound": "hbm", "achieved": gbs, "peak": peak, "unit": "GB/s", "frac": gbs / peak,
                                                         "per": "GPU"}}
            if sort:
                if any_hit:
                    same = bool(((hits[:, 3].view(torch.int32) >= 0) == (hits2[:, 3].view(torch.int32) >= 0)).all().item())
                else:
                    same = bool(torch.equal(hits.view(torch.int32), hits2.view(torch.int32)))
                cast["parity"] = {"ok": same, "against": "the unsorted cast of the same rays (bit-equal hits)"}
            res["casts"].append(cast)
    # e2e: host buffers through frt_trace_rays (copies inside), a bounded block per rank
    m = min(n, 1 << 24)
    pr, ph = api.Pinned(m, api.RAY_DTYPE), api.Pinned(m, api.HIT_DTYPE)
    pr.array.view(np.float32).reshape(m, 6)[:] = sets["closest"][0][:m].cpu().numpy()
    rig.ctx.trace(pr.array, hits=ph.array)
    rig.barrier()
    t0 = time.perf_counter()
    _, dms = rig.ctx.trace(pr.array, hits=ph.array)
    wall = rig.max_over_ranks(time.perf_counter() - t0)
    res["e2e"] = {"value": m * rig.world / wall / 1e6, "unit": "Mrays/s", "rays_per_rank": m, "seconds": wall, "device_ms": dms,
                  "h2d_bytes": m * 24, "d2h_bytes": m * 16, "call": "frt_trace_rays, pinned host rays in / hits out, closest hit, unsorted"}
    pr.close(); ph.close()
    # what ncu measured on this kernel (another launch size: 2^22 rays; per-ray figures carry over): real DRAM and L2 bytes
    m, src = ncu_metrics()
    if m:
        ncu = {}
        for key, label in (("k_trace", "closest"), ("k_trace_any", "shadow")):
            k = m.get(key) or {}
            if k.get("dram_bytes") and k.get("duration_s"):
                rays_captured = 1 << 22
                ncu[label] = {"dram_bytes_per_ray": k["dram_bytes"] / rays_captured, "l2_bytes_per_ray": (k.get("lts_bytes") or 0) / rays_captured,
                              "dram_gbs": k["dram_bytes"] / k["duration_s"] / 1e9, "dram_frac_of_peak": k["dram_bytes"] / k["duration_s"] / 1e9 / peak,
                              "issue_active_frac": (k.get("issue_active_pct") or 0) / 100.0, "active_lanes_per_inst": k.get("lanes_per_inst"),
                              "capture": k.get("capture")}
        res["ncu"] = dict(ncu, source=src)
    res["clocks"] = sampler.finish()
    res["sm_mhz_per_rank"] = rig.per_rank(res["clocks"].get("sm_mhz") or 0.0)
    scene.close()
    return res


def e2e_pass(scene, steps, n_gpus, spp):
    """The call a user of the reference makes: renderer<gpu_path_integrator>::Render(scene, film)
    (frth_render) with HOST scene in and HOST film out — scene flatten + upload, render, multi-GPU film sum,
    device tone map, device->host film copy and the viewer's fout_image / out_image fill all inside."""
    scene.render(64, 64, 4, n_gpus=n_gpus)  # CUDA context / module load warm-up
    scene.render(NX, NY, 1, n_gpus=n_gpus)  # film-sized buffers (pinned staging, the viewer) exist from here on
    t0 = time.perf_counter()
    for _ in range(steps):
        img, st, secs = scene.render(NX, NY, spp, seed=1, n_gpus=n_gpus)
    dt = time.perf_counter() - t0
    return steps * NX * NY * spp / dt / 1e6, float(img.mean()), dt / steps, st["render_ms"]


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
    rig = Rig()
    rank, world = rig.rank, rig.world
    torch, dist = rig.torch, rig.dist
    peaks = {}
    try:
        peaks = json.loads((REPO / "MEASURED_PEAKS.json").read_text())
    except Exception:
        pass
    spp, steps = args.spp, args.steps
    res, scene, (s0, s1) = run_headline(rig, args, spp)
    total_s = res["total_ms"] * 1e-3
    value = steps * NX * NY * spp / total_s / 1e6
    roof = cpu = e2e = None
    if rank == 0 and not args.no_roofline:
        roof = roofline_pass(rig, (0, max(1, spp // world)), peaks, res["clocks"])
    configs = None
    if not args.no_configs:
        configs = {}
        want_cpu = world == 1 and not args.no_cpu_baseline
        configs["config1_cornell_512x512x16"] = config1(rig, want_cpu)
        configs["config3_veach_768x512x256"] = config_render(rig, "veach", 768, 512, 256, fixture="stat_veach_768x512x256_block4")
        configs["config3_veach_1920x1080x256"] = config_render(rig, "veach", 1920, 1080, 256)
        configs["config4_teapot_1920x1080x256"] = config_render(rig, "teapot", 1920, 1080, 256, fixture="stat_teapot_480x270x256")
        configs["config5_soup"] = config_soup(rig, int(args.soup_tris), args.soup_rays, peaks)
    rig.ctx.close()
    # The end-to-end leg runs renderer<gpu_path_integrator> with n_gpus = world inside rank 0's process; the other
    # ranks must leave their GPUs idle meanwhile, so they wait on a CPU (gloo) barrier: an NCCL barrier spins a
    # kernel on the waiting GPU.
    cpu_group = dist.new_group(backend="gloo") if world > 1 else None
    if world > 1:
        torch.cuda.synchronize()
        dist.barrier(group=cpu_group)
    if rank == 0 and not args.no_e2e:
        e2e_val, e2e_mean, e2e_s, e2e_dev_ms = e2e_pass(scene, max(1, min(steps, 3)), world, spp)
        e2e = {"value": e2e_val, "unit": "Msamples/s", "h2d_bytes_per_step": scene_bytes(scene),
               "d2h_bytes_per_step": NX * NY * 3 * 4 + NX * NY * 3 + 128 * world, "seconds_per_step": e2e_s,
               "device_ms_per_step": e2e_dev_ms, "mean_radiance": e2e_mean,
               "call": "renderer<gpu_path_integrator>::Render via frth_render (host scene in; host film out: fout_image, "
                       "out_image tone-mapped on the device, float32 mean)"}
        try:
            ref_mean = float(np.load(GOLDEN / "stat_cornell_1920x1080x256_block8.npz")["mean"])
            if SCENE == "cornell" and spp == SPP:
                e2e["parity"] = {"ok": bool(abs(e2e_mean - ref_mean) <= 5e-3 * ref_mean), "mean_rel_diff": abs(e2e_mean - ref_mean) / ref_mean,
                                 "against": f"tests/golden/stat_cornell_1920x1080x256_block8.npz (reference path::Li, mt19937, mean {ref_mean:.6f})"}
        except Exception:
            pass
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
                       "pool": "CTA-resident (shared memory, one persistent kernel per tile)",
                       "timing": "CUDA events on the launching stream per step, summed, max over ranks"},
            "clocks": res["clocks"], "e2e": e2e, "gpu_launches": res["launches"],
            "roofline": roof, "cpu_baseline": cpu, "configs": configs,
        }
        emit(line)
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
