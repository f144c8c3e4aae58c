"""Statistical-parity fixtures from the reference's OWN object code with its as-shipped std::mt19937 sampler
(oracle/_ref/frt_ref, compiled from /root/reference by oracle/Makefile).  Run in the build container:

    python scripts/make_golden_stat.py [--only NAME]

The reference scales 1.3x on 8 taskflow threads (global atomics), so the long renders are run as one
single-thread process per row band (the sampler is seeded per row, path.cpp:122, so bands are independent) and
stitched.  Config 1 (512x512x16, the one BASELINE config the reference can run verbatim) is ALSO run as shipped — one
process, taskflow on every core — and that run's seconds are what the fixture records as the reference timing.

Fixtures:
  stat_cornell_1920x1080x256_block8.npz  BASELINE config 2 at its own resolution: 8x8 block means (135x240x3)
  stat_cornell_512x512x1024.npz          config 1 converged (full resolution)
  stat_cornell_512x512x16.npz            config 1 itself + timing of the as-shipped threaded run
  stat_cornell_256x256x1024.npz          the 256^2 statistical test at the 0.5 % bar
  stat_veach_768x512x256_block4.npz      config 3 film size (mi.xml:144-145), 4x4 block means
  stat_teapot_480x270x256.npz            config 4 stand-in, reduced
"""
from __future__ import annotations

import argparse
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO))
sys.path.insert(0, str(REPO / "tests"))

import refutil  # noqa: E402

OUT = REPO / "tests" / "golden"


def banded(scene, nx, ny, spp, seed, procs=None):
    procs = procs or os.cpu_count() or 1
    bands = [(ny * i // procs, ny * (i + 1) // procs) for i in range(procs)]
    bands = [b for b in bands if b[1] > b[0]]
    t0 = time.time()

    def one(b):
        img, info = refutil.ref_render(scene, nx, ny, spp, ctr=False, seed=seed, threads=1, rows=b)
        return b, img, info

    img = np.zeros((ny, nx, 3), np.float32)
    tot = {"closest_rays": 0, "shadow_rays": 0, "nonfinite": 0, "cpu_seconds": 0.0}
    with ThreadPoolExecutor(len(bands)) as ex:
        for (y0, y1), part, info in ex.map(one, bands):
            img[y0:y1] = part[y0:y1]
            tot["closest_rays"] += info["closest_rays"]; tot["shadow_rays"] += info["shadow_rays"]
            tot["nonfinite"] += info["nonfinite"]; tot["cpu_seconds"] += info["seconds"]
    tot["wall_seconds"] = time.time() - t0
    return img, tot


def save(name, img, tot, seed, **extra):
    np.savez_compressed(OUT / name, image=img.astype(np.float32), seed=seed, mean=np.float64(extra.pop("mean")),
                        closest_rays=tot["closest_rays"], shadow_rays=tot["shadow_rays"], nonfinite=tot["nonfinite"],
                        cpu_seconds=tot["cpu_seconds"], **extra)
    print(f"{name}: mean {float(img.mean()):.6f}, {tot['cpu_seconds']:.0f} CPU-s, {tot['wall_seconds']:.0f} s wall", flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="")
    args = ap.parse_args()
    if not refutil.have_ref():
        raise SystemExit("oracle/_ref is missing: run `make -C oracle` where /root/reference exists")

    def want(n):
        return not args.only or args.only in n

    if want("cornell_512x16"):
        # config 1 as shipped: ONE process, taskflow on every core (main.cpp:222-252 scene, path::Li, full rows)
        cores = os.cpu_count() or 1
        img, info = refutil.ref_render("cornell", 512, 512, 16, ctr=False, seed=5, threads=cores)
        _, info1 = refutil.ref_render("cornell", 512, 512, 16, ctr=False, seed=5, threads=1)
        np.savez_compressed(OUT / "stat_cornell_512x512x16.npz", image=img.astype(np.float32), seed=5, mean=np.float64(img.mean()),
                            closest_rays=info["closest_rays"], shadow_rays=info["shadow_rays"], nonfinite=info["nonfinite"],
                            seconds=info["seconds"], threads=cores, seconds_1thread=info1["seconds"],
                            samples_per_s=info["samples_per_s"], mrays_per_s=info["mrays_per_s"])
        print(f"stat_cornell_512x512x16: {info['seconds']:.2f} s on {cores} threads, {info1['seconds']:.2f} s on 1, mean {img.mean():.6f}", flush=True)
    if want("cornell_256"):
        img, tot = banded("cornell", 256, 256, 1024, seed=5)
        save("stat_cornell_256x256x1024.npz", img, tot, 5, mean=img.mean())
    if want("cornell_512x1024"):
        img, tot = banded("cornell", 512, 512, 1024, seed=5)
        save("stat_cornell_512x512x1024.npz", img, tot, 5, mean=img.mean())
    if want("cornell_1080"):
        img, tot = banded("cornell", 1920, 1080, 256, seed=5)
        save("stat_cornell_1920x1080x256_block8.npz", refutil.block_mean(img, 8), tot, 5, mean=img.mean(), block=8)
    if want("veach_768"):
        img, tot = banded("veach", 768, 512, 256, seed=5)
        save("stat_veach_768x512x256_block4.npz", refutil.block_mean(img, 4), tot, 5, mean=img.mean(), block=4)
    if want("teapot_480"):
        img, tot = banded("teapot", 480, 270, 256, seed=5)
        save("stat_teapot_480x270x256.npz", img, tot, 5, mean=img.mean())


if __name__ == "__main__":
    main()
