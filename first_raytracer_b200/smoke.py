"""smoke(): one small invocation of the hot path on cuda:0, checked against the oracle.

Renders Cornell 64x64 @ 8 spp through renderer<gpu_path_integrator> (C++ host layer → C ABI →
CUDA wavefront kernels) and casts 4096 camera rays through frt_trace_rays, then compares
  * the image with the reference's own path::Li fed the same counter-based random numbers
    (oracle/_ref/frt_ref_ctr, compiled from the reference sources; test infrastructure), and
  * the hits with the C restatement oracle/libfrt_oracle.so (float64 brute force).
The oracle is only the checker here; nothing below runs when the CUDA library is missing.
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parent.parent


def run() -> None:
    sys.path.insert(0, str(REPO))
    sys.path.insert(0, str(REPO / "tests"))
    from first_raytracer_b200 import api

    scene = api.HostScene("cornell", 1.0, 4)
    nx = ny = 64
    spp = 8
    img, st, secs = scene.render(nx, ny, spp, seed=7)
    assert np.isfinite(img).all() and img.mean() > 0.01, "GPU render is empty"
    assert st["samples"] == nx * ny * spp and st["kernel_launches"] > 0
    print(f"smoke: cornell {nx}x{ny}x{spp}: mean {img.mean():.5f}, {st['closest_rays']}+{st['shadow_rays']} rays, "
          f"{st['kernel_launches']} launches, {st['render_ms']:.2f} ms device")

    ctx = api.Context()
    ctx.upload(scene)
    rays = scene.camera_rays(64, 64)
    hits, ms = ctx.trace(rays)
    assert (hits["prim_id"] >= 0).mean() > 0.5

    import refutil
    checked = []
    try:
        import oracle_c
        ref = oracle_c.Oracle(scene).trace(rays)
        agree = (ref["prim"] == hits["prim_id"]).mean()
        assert agree >= 0.999, f"closest-hit ids differ from the C oracle: {agree}"
        checked.append(f"C oracle id agreement {agree:.4f}")
        oimg = oracle_c.Oracle(scene).render(nx, ny, spp, seed=7)
        d = abs(oimg.mean() - img.mean()) / oimg.mean()
        assert d < 1e-2, f"mean radiance differs from the C oracle by {d:.3%}"
        checked.append(f"C oracle mean diff {d:.2e}")
    except ImportError:
        pass
    if refutil.have_ref():
        rimg, _ = refutil.ref_render("cornell", nx, ny, spp, ctr=True, seed=7)
        d = abs(rimg.mean() - img.mean()) / rimg.mean()
        assert d < 1e-2, f"mean radiance differs from the compiled reference by {d:.3%}"
        checked.append(f"reference (same RNG) mean diff {d:.2e}")
    ctx.close()
    assert checked, "no oracle available to check against"
    print("smoke: ok — " + "; ".join(checked))
