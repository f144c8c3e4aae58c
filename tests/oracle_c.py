"""ctypes wrapper of the C restatement oracle/libfrt_oracle.so (test infrastructure only)."""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import numpy as np

from first_raytracer_b200 import api

REPO = Path(__file__).resolve().parent.parent
_cache = {}


class frt_oracle_stats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("samples", "closest_rays", "shadow_rays", "nonfinite", "node_visits", "tri_tests")]


def lib(f32: bool = False):
    key = "f32" if f32 else "f64"
    if key not in _cache:
        p = REPO / "oracle" / ("libfrt_oracle_f32.so" if f32 else "libfrt_oracle.so")
        if not p.exists():
            raise ImportError(f"{p} is missing (make -C oracle)")
        L = C.CDLL(str(p))
        vp = C.c_void_p
        L.frt_oracle_trace.restype = C.c_int
        L.frt_oracle_trace.argtypes = [vp, vp, C.c_size_t, C.c_int, C.c_int, vp, vp, vp, vp]
        L.frt_oracle_render.restype = C.c_int
        L.frt_oracle_render.argtypes = [vp, C.POINTER(api.frt_tile), C.c_int, vp, C.POINTER(frt_oracle_stats)]
        L.frt_oracle_bsdf.restype = C.c_int
        L.frt_oracle_bsdf.argtypes = [vp, vp, vp, vp, vp, vp]
        L.frt_oracle_prim_t.restype = C.c_int
        L.frt_oracle_prim_t.argtypes = [vp, vp, vp, C.c_size_t, vp]
        L.frt_oracle_is_f32.restype = C.c_int
        assert L.frt_oracle_is_f32() == int(f32)
        _cache[key] = L
    return _cache[key]


class Oracle:
    """The restated hot path over a flattened scene description (api.HostScene or a raw desc)."""

    def __init__(self, scene, f32: bool = False, use_bvh: bool | None = None):
        self.scene = scene
        self.desc = scene.desc if hasattr(scene, "desc") else scene
        self.lib = lib(f32)
        n_prims = self.desc.n_tris + self.desc.n_spheres
        self.use_bvh = (n_prims > 64) if use_bvh is None else use_bvh

    def trace(self, rays: np.ndarray, any_hit: bool = False, use_bvh: bool | None = None):
        rays = np.ascontiguousarray(rays)
        n = rays.shape[0]
        prim = np.zeros(n, np.int32)
        t, u, v = np.zeros(n), np.zeros(n), np.zeros(n)
        rc = self.lib.frt_oracle_trace(C.addressof(self.desc), rays.ctypes.data, n, int(any_hit),
                                       int(self.use_bvh if use_bvh is None else use_bvh),
                                       prim.ctypes.data, t.ctypes.data, u.ctypes.data, v.ctypes.data)
        assert rc == 0
        return {"prim": prim, "t": t, "u": u, "v": v}

    def prim_t(self, rays: np.ndarray, prim_id: np.ndarray) -> np.ndarray:
        """t at which each ray hits the ONE given primitive (NaN if it does not)."""
        rays = np.ascontiguousarray(rays)
        ids = np.ascontiguousarray(prim_id, dtype=np.int32)
        t = np.zeros(rays.shape[0])
        rc = self.lib.frt_oracle_prim_t(C.addressof(self.desc), rays.ctypes.data, ids.ctypes.data, rays.shape[0], t.ctypes.data)
        assert rc == 0
        return t

    def id_agreement(self, rays, prim, ref_prim, ref_t, tie_rel=1e-6):
        """(strict, tie_adjusted) fraction of rays whose primitive id equals the reference's.  A
        mismatch counts as a tie when the other primitive, tested alone in float64, is hit at the
        reference's t within tie_rel: coincident surfaces, where the reference's own winner is an
        artefact of traversal order and last-bit rounding."""
        prim, ref_prim = np.asarray(prim), np.asarray(ref_prim)
        same = prim == ref_prim
        bad = np.flatnonzero(~same & (prim >= 0) & (ref_prim >= 0))
        tie = np.zeros_like(same)
        if bad.size:
            t_other = self.prim_t(rays[bad], prim[bad])
            tie[bad] = np.abs(t_other - ref_t[bad]) <= tie_rel * np.abs(ref_t[bad])
        return float(same.mean()), float((same | tie).mean())

    def render(self, nx, ny, spp, seed=1, spp_begin=0, max_depth=34, rect=None, return_stats=False, integrator="path"):
        x0, y0, w, h = rect if rect else (0, 0, nx, ny)
        tile = api.frt_tile(nx, ny, x0, y0, w, h, spp_begin, spp_begin + spp, seed, max_depth, api.INTEGRATORS[integrator], 0)
        img = np.zeros((ny, nx, 3), np.float64)
        st = frt_oracle_stats()
        rc = self.lib.frt_oracle_render(C.addressof(self.desc), C.byref(tile), int(self.use_bvh), img.ctypes.data, C.byref(st))
        assert rc == 0
        img /= spp
        if return_stats:
            return img, {n: getattr(st, n) for n, _ in st._fields_}
        return img

    @staticmethod
    def bsdf(material: np.ndarray, n, wi, s, wo_probe, f32: bool = False):
        """material: one element of api.MATERIAL_DTYPE → 12 doubles (see frt_oracle.h)."""
        m = np.ascontiguousarray(material)
        out = np.zeros(12)
        a = [np.ascontiguousarray(x, dtype=np.float64) for x in (n, wi, s, wo_probe)]
        rc = lib(f32).frt_oracle_bsdf(m.ctypes.data, a[0].ctypes.data, a[1].ctypes.data, a[2].ctypes.data,
                                      a[3].ctypes.data, out.ctypes.data)
        assert rc == 0
        return out
