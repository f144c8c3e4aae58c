"""Test helpers: run the compiled reference (oracle/_ref, built from /root/reference by
oracle/Makefile) and the C restatement (oracle/libfrt_oracle.so).  Test infrastructure only."""
from __future__ import annotations

import json
import os
import subprocess
import tempfile
from pathlib import Path

import numpy as np

REPO = Path(__file__).resolve().parent.parent
REF_BIN = REPO / "oracle" / "_ref" / "frt_ref"
REF_CTR_BIN = REPO / "oracle" / "_ref" / "frt_ref_ctr"
ASSETS = REPO / "assets"
GOLDEN = REPO / "tests" / "golden"


def have_ref() -> bool:
    return REF_BIN.exists() and REF_CTR_BIN.exists()


def _run(cmd):
    env = dict(os.environ, FRT_ASSETS=str(ASSETS))
    out = subprocess.run([str(c) for c in cmd], capture_output=True, text=True, env=env, timeout=3600)
    if out.returncode != 0:
        raise RuntimeError(f"{cmd[0]} failed ({out.returncode}): {out.stderr[-2000:]}")
    return json.loads(out.stdout.strip().splitlines()[-1])


def read_pfm(path) -> np.ndarray:
    with open(path, "rb") as f:
        assert f.readline().strip() == b"PF"
        w, h = map(int, f.readline().split())
        f.readline()
        return np.frombuffer(f.read(), dtype="<f4").reshape(h, w, 3).copy()


def write_pfm(path, img: np.ndarray):
    h, w, _ = img.shape
    with open(path, "wb") as f:
        f.write(f"PF\n{w} {h}\n-1\n".encode())
        f.write(np.ascontiguousarray(img, dtype="<f4").tobytes())


def ref_trace(scene: str, aspect: float, rays: np.ndarray, shadow: bool = False):
    """scene->world->hit on the reference's own BVH/triangle/sphere code → dict(prim,t,u,v) + info."""
    n = rays.shape[0]
    with tempfile.TemporaryDirectory() as td:
        rp, op = Path(td) / "rays.f32", Path(td) / "out.bin"
        np.ascontiguousarray(rays).view(np.float32).reshape(n, 6).tofile(rp)
        info = _run([REF_BIN, "trace", scene, aspect, "shadow" if shadow else "closest", rp, op])
        raw = op.read_bytes()
    prim = np.frombuffer(raw, dtype="<i4", count=n)
    rest = np.frombuffer(raw, dtype="<f8", offset=4 * n).reshape(3, n)
    return {"prim": prim.copy(), "t": rest[0].copy(), "u": rest[1].copy(), "v": rest[2].copy()}, info


def ref_render(scene: str, nx: int, ny: int, spp: int, ctr: bool = False, seed: int = 1, threads: int | None = None,
               rows=None, ystep: int = 1, integrator: str = "path"):
    """path::Li / ao::Li / normals_renderer::Li over the full frame (or a row band) → (image [ny,nx,3] row 0 = bottom, info)."""
    with tempfile.TemporaryDirectory() as td:
        out = Path(td) / "img.pfm"
        cmd = [REF_CTR_BIN if ctr else REF_BIN, "render", scene, nx, ny, spp, out, "--seed", seed]
        if threads:
            cmd += ["--threads", threads]
        if rows:
            cmd += ["--rows", rows[0], rows[1]]
        if ystep != 1:
            cmd += ["--ystep", ystep]
        if integrator != "path":
            cmd += ["--integrator", integrator]
        info = _run(cmd)
        img = read_pfm(out)
    return img, info


def ref_dump(scene: str, aspect: float):
    with tempfile.TemporaryDirectory() as td:
        out = Path(td) / "scene.bin"
        info = _run([REF_BIN, "dump", scene, aspect, out])
        raw = np.fromfile(out, dtype="<f8")
    assert raw[0] == 3.0
    cam = raw[1:23]
    n_prims, n_lights = int(raw[23]), int(raw[24])
    rec = raw[25:].reshape(n_prims, 1 + 18 + 2 + 1 + 12)
    return {"camera": cam, "n_lights": n_lights, "kind": rec[:, 0], "geom": rec[:, 1:19].reshape(n_prims, 6, 3),
            "inv_area": rec[:, 19], "geo_normals": rec[:, 20], "mat_kind": rec[:, 21], "mat_p": rec[:, 22:34]}, info


def rel_mse(a: np.ndarray, b: np.ndarray) -> float:
    """mean((a-b)^2 / (b^2 + 1e-4))"""
    a = a.astype(np.float64)
    b = b.astype(np.float64)
    return float(np.mean((a - b) ** 2 / (b ** 2 + 1e-4)))


def block_mean(img: np.ndarray, k: int) -> np.ndarray:
    h, w, c = img.shape
    return img[: h - h % k, : w - w % k].reshape(h // k, k, w // k, k, c).mean(axis=(1, 3))


def random_rays(n: int, lo, hi, seed: int) -> np.ndarray:
    """Incoherent rays: origins uniform in the box [lo,hi], directions uniform on the sphere scaled to
    a length in [0.5, 10]."""
    from first_raytracer_b200.api import RAY_DTYPE
    rng = np.random.default_rng(seed)
    lo, hi = np.asarray(lo, np.float64), np.asarray(hi, np.float64)
    o = lo + rng.random((n, 3)) * (hi - lo)
    d = rng.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d *= rng.uniform(0.5, 10.0, size=(n, 1))
    rays = np.zeros(n, dtype=RAY_DTYPE)
    rays["o"] = o.astype(np.float32)
    rays["d"] = d.astype(np.float32)
    return rays


def segment_rays(n: int, lo, hi, seed: int) -> np.ndarray:
    """Shadow segments p→q between two uniform points of the box."""
    from first_raytracer_b200.api import RAY_DTYPE
    rng = np.random.default_rng(seed)
    lo, hi = np.asarray(lo, np.float64), np.asarray(hi, np.float64)
    p = lo + rng.random((n, 3)) * (hi - lo)
    q = lo + rng.random((n, 3)) * (hi - lo)
    rays = np.zeros(n, dtype=RAY_DTYPE)
    rays["o"] = p.astype(np.float32)
    rays["d"] = (q - p).astype(np.float32)
    return rays


DROPIN_BIN = REPO / "oracle" / "_ref" / "frt_dropin"


def dropin_flatten(scene: str, aspect: float, max_leaf: int) -> dict:
    """oracle/_ref/frt_dropin flatten: the product's flattener compiled against the REFERENCE's headers, walking the
    reference's own Scene → raw bytes of each frt_scene_desc array."""
    with tempfile.TemporaryDirectory() as td:
        out = Path(td) / "flat.bin"
        _run([DROPIN_BIN, "flatten", scene, aspect, max_leaf, out])
        raw = out.read_bytes()
    head = np.frombuffer(raw, "<u4", 16)
    n_nodes, n_tris, n_sph, n_mat, n_lights, has_n, has_env, env_nx, env_ny, n_tex, n_texels, has_uv = (int(x) for x in head[:12])
    pos = 64
    parts = {}

    def take(name, nbytes):
        nonlocal pos
        parts[name] = raw[pos:pos + nbytes]
        pos += nbytes

    take("nodes", 128 * n_nodes); take("tris", 48 * n_tris)
    take("tri_normals", 48 * n_tris if has_n else 0)
    take("spheres", 32 * n_sph); take("materials", 64 * n_mat); take("lights", 128 * n_lights)
    take("camera", 22 * 4); take("env_color", 12); take("ao_t_max", 4)
    take("textures", 32 * n_tex); take("tex_texels", 12 * n_texels); take("tri_uv", 24 * n_tris if has_uv else 0)
    assert pos == len(raw)
    parts["has_env_image"] = bool(has_env)
    return parts


def dropin_render(scene: str, nx: int, ny: int, spp: int, seed: int = 1, gpus: int = 1, integrator: str = "path"):
    """oracle/_ref/frt_dropin gpurender: the reference's own renderer<> (integrator.h:11-47) instantiated with
    gpu_path_integrator, on the reference's Scene → (mean image, info)."""
    with tempfile.TemporaryDirectory() as td:
        out = Path(td) / "img.pfm"
        info = _run([DROPIN_BIN, "gpurender", scene, nx, ny, spp, out, "--seed", seed, "--gpus", gpus, "--integrator", integrator])
        img = read_pfm(out)
    return img, info
