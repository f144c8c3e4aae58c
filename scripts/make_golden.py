"""Generates tests/golden/*.npz from the reference's OWN object code (oracle/_ref/frt_ref*, compiled
from /root/reference by oracle/Makefile).  Run in the build container, where /root/reference
exists; the fixtures are committed so that the oracle stays pinned on machines without it.

    python scripts/make_golden.py

Fixtures:
  trace_<scene>.npz      F1/F2: rays + closest-hit (prim id, t, u, v) and segments + occluded flag
  render_<scene>_WxHxS   F3: path::Li images with the counter-based sampler (same numbers as the GPU)
  bsdf_vectors.npz       F4: scatter / generate / value / eval_bsdf of every material class
"""
from __future__ import annotations

import subprocess
import sys
import tempfile
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO))
sys.path.insert(0, str(REPO / "tests"))

import refutil  # noqa: E402
from first_raytracer_b200 import api  # noqa: E402

OUT = REPO / "tests" / "golden"
SEED = 7


def flat(rays):
    return np.concatenate([rays["o"], rays["d"]], axis=1).astype(np.float32)


def trace_fixture(name, aspect, n_cam=48, n_rand=6000, n_seg=4000):
    scene = api.HostScene(name, aspect, 4)
    info = scene.info
    rays = np.concatenate([scene.camera_rays(n_cam, n_cam),
                           refutil.random_rays(n_rand, info["bounds_lo"], info["bounds_hi"], seed=101)])
    ref, _ = refutil.ref_trace(name, aspect, rays)
    seg = refutil.segment_rays(n_seg, info["bounds_lo"], info["bounds_hi"], seed=102)
    sref, _ = refutil.ref_trace(name, aspect, seg, shadow=True)
    np.savez_compressed(OUT / f"trace_{name}.npz", rays=flat(rays), prim=ref["prim"], t=ref["t"], u=ref["u"], v=ref["v"],
                        segments=flat(seg), occluded=sref["prim"] >= 0)
    print(f"trace_{name}: {rays.shape[0]} rays, hit fraction {np.mean(ref['prim'] >= 0):.3f}, "
          f"{seg.shape[0]} segments, occluded {np.mean(sref['prim'] >= 0):.3f}")


def render_fixture(name, nx, ny, spp, integrator="path"):
    """integrator: path::Li, ao::Li (ao.cpp:4-28) or normals_renderer::Li (debug_renderer.h:8-17); a scene name
    ending in +sky is the scene under a constant sky-blue environment, in +hdr under the HDR lat-long image
    assets/data/test.hdr read by the reference's own stb loader (ref_main.cpp build_scene)."""
    img, info = refutil.ref_render(name, nx, ny, spp, ctr=True, seed=SEED, integrator=integrator)
    tag = "render" if integrator == "path" else f"render_{integrator}"
    np.savez_compressed(OUT / f"{tag}_{name}_{nx}x{ny}x{spp}.npz", image=img.astype(np.float32), seed=SEED,
                        closest_rays=info["closest_rays"], shadow_rays=info["shadow_rays"], nonfinite=info["nonfinite"])
    print(f"{tag}_{name}_{nx}x{ny}x{spp}: mean {img.mean():.6f}, rays {info['closest_rays']}+{info['shadow_rays']}")


def unit(v):
    return v / np.linalg.norm(v, axis=-1, keepdims=True)


def bsdf_fixture(n_per=600):
    rng = np.random.default_rng(5)
    cu = [1.65746, 0.880369, 0.521229]
    ck = [9.22387, 6.26952, 4.837]
    mats = [(0, [.7, .6, .5] + [0] * 8), (1, [.3, .2, .1, .5, .6, .7, 20] + [0] * 4), (1, [.1, .1, .1, .9, .9, .9, 1024] + [0] * 4),
            (2, [.9, .9, .9, 0, 0, 0, 1.5] + [0] * 4), (2, [.3, .3, .3, 0, 0, 0, 2.5] + [0] * 4),
            (3, [1, 1, 1] + cu + [0.005] + ck + [0]), (3, [1, 1, 1] + cu + [0.05] + ck + [0]), (3, [1, 1, 1] + cu + [0.1] + ck + [0]),
            (3, [1, 1, 1] + cu + [0.1] + ck + [1]), (3, [.8, .9, 1] + cu + [0.3] + ck + [1]), (4, [.8, .7, .6] + [0] * 8)]
    rec = np.zeros((len(mats) * n_per, 24))
    for k, (kind, p) in enumerate(mats):
        r = rec[k * n_per:(k + 1) * n_per]
        r[:, 0] = kind
        r[:, 1:12] = np.asarray(p, np.float32)  # what a frt_material can hold
        n = unit(rng.normal(size=(n_per, 3)))
        wi = unit(rng.normal(size=(n_per, 3)))
        flip = (np.sum(n * wi, axis=1) < 0) & (rng.random(n_per) < 0.9)
        wi[flip] *= -1
        r[:, 12:15], r[:, 15:18] = n, wi
        r[:, 18:21] = np.floor(rng.random((n_per, 3)) * 2 ** 24) / 2 ** 24  # 24-bit uniforms like the RNG
        r[:, 21:24] = unit(n + 0.7 * unit(rng.normal(size=(n_per, 3))))
    with tempfile.TemporaryDirectory() as td:
        rec.tofile(td + "/in")
        subprocess.run([str(refutil.REF_BIN), "bsdf", td + "/in", td + "/out"], check=True, capture_output=True)
        out = np.fromfile(td + "/out").reshape(-1, 12)
    np.savez_compressed(OUT / "bsdf_vectors.npz", records=rec, outputs=out)
    print(f"bsdf_vectors: {rec.shape[0]} records, finite {np.isfinite(out).all(axis=1).mean():.4f}")


OTHER_INTEGRATOR_CASES = [("teapot", 48, 27, 2, "ao"), ("cornell_glossy+sky", 32, 32, 2, "ao"), ("veach+sky", 48, 32, 2, "ao"),
                          ("cornell_sphere+sky", 32, 32, 2, "ao"), ("cornell", 32, 32, 1, "normals"),
                          ("teapot", 48, 27, 1, "normals"), ("cornell_sphere", 32, 32, 1, "normals"),
                          ("veach+sky", 48, 32, 2, "path"), ("teapot", 160, 90, 4, "ao"),
                          # ass_1_renderer (assignment_1.cpp:4-38); cornell_sphere is left out: modified_phong::eval_bsdf
                          # with the un-normalised to_light overflows float there (the reference's film holds inf)
                          ("cornell", 32, 32, 2, "direct"), ("cornell_glossy", 32, 32, 2, "direct"),
                          ("teapot", 48, 27, 2, "direct"), ("teapot+sky", 48, 27, 1, "direct"),
                          ("cornell", 128, 128, 4, "direct"),
                          # environment_map over the HDR image of main.cpp:170 (material.h:209-232, texture.h:60-86)
                          ("teapot+hdr", 48, 27, 2, "path"), ("veach+hdr", 48, 32, 2, "path"),
                          ("teapot+hdr", 160, 90, 4, "path")]


# Round 2: the branches no shipped scene reaches — Beckmann rough_conductor (pdf.h:280-403), `metal`
# (material.h:111-130), the thin-lens camera (camera.h:10-35, util.h:21-41) — and surface textures (texture.h:30-112)
# on the reduced random_scene of ref_main.cpp "textured" (ao is left out there: constant_pdf::generate throws, pdf.h:197).
ROUND2_TRACE = [("textured", 1.5), ("random_scene", 1.5)]
ROUND2_RENDER = [("veach_beckmann", 48, 32, 4, "path"), ("veach_beckmann", 192, 128, 16, "path"),
                 ("cornell_metal", 32, 32, 2, "path"), ("cornell_metal", 96, 96, 8, "path"),
                 ("cornell_lens", 32, 32, 2, "path"), ("cornell_lens", 96, 96, 8, "path"),
                 ("textured", 48, 32, 2, "path"), ("textured", 192, 128, 8, "path"), ("textured+hdr", 48, 32, 2, "path"),
                 ("textured", 96, 64, 2, "direct"), ("textured", 96, 64, 1, "normals"),
                 ("veach_beckmann+sky", 48, 32, 2, "ao"),
                 # furnace_test_scene (main.cpp:177-221): dielectric teapot, NO lights, image environment
                 ("furnace+hdr", 48, 27, 2, "path"), ("furnace+hdr", 160, 90, 4, "path"),
                 # cornell_box_ao (main.cpp:254-280): Cornell with Assimp's smooth vertex normals, white environment
                 ("cornell_ao", 32, 32, 2, "ao"), ("cornell_ao", 96, 96, 8, "ao"), ("cornell_ao", 32, 32, 2, "path"),
                 ("cornell_ao", 32, 32, 1, "normals"),
                 # random_scene (main.cpp:55-113), the std::mt19937(1) instance: 485 spheres, thin lens
                 ("random_scene", 48, 32, 2, "path"), ("random_scene", 192, 128, 8, "path"), ("random_scene+hdr", 48, 32, 2, "path"),
                 ("random_scene", 96, 64, 1, "normals")]


def main():
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("--round2-only", action="store_true", help="only the fixtures added in round 2")
    args = ap.parse_args()
    if not refutil.have_ref():
        raise SystemExit("oracle/_ref is missing: run `make -C oracle` where /root/reference exists")
    OUT.mkdir(parents=True, exist_ok=True)
    for name, aspect in ROUND2_TRACE:
        trace_fixture(name, aspect)
    for name, nx, ny, spp, integ in ROUND2_RENDER:
        render_fixture(name, nx, ny, spp, integ)
    if args.round2_only:
        return
    for name, aspect in [("cornell", 1.0), ("veach", 1.5), ("teapot", 16 / 9), ("cornell_sphere", 1.0)]:
        trace_fixture(name, aspect)
    for name, nx, ny, spp in [("cornell", 48, 48, 4), ("veach", 48, 32, 4), ("cornell_sphere", 32, 32, 2),
                              ("cornell_glossy", 32, 32, 2), ("teapot", 48, 27, 2),
                              # larger ones for the GPU parity tests (statistical bars need pixels x spp)
                              ("cornell", 128, 128, 16), ("veach", 192, 128, 16)]:
        render_fixture(name, nx, ny, spp)
    # the integrators next to `path` and the constant-environment term of path::Li
    for name, nx, ny, spp, integ in OTHER_INTEGRATOR_CASES:
        render_fixture(name, nx, ny, spp, integ)
    bsdf_fixture()


if __name__ == "__main__":
    main()
