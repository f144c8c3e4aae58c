"""ctypes binding of the two C ABIs (include/frt.h → libfrt.so, include/frt_host.h → libfrt_host.so).

Python is only the test / bench harness language here: the product is the CUDA core and the C++
host layer.  Nothing in this module computes a hit or a pixel; if the shared libraries are
missing it raises (there is no Python or CPU fallback).
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

PKG_DIR = Path(__file__).resolve().parent
REPO_DIR = PKG_DIR.parent

FRT_OK = 0
FRTH_CANCELLED = 1
FRT_FLAG_COUNTERS = 1
FRT_TRACE_ANY_HIT = 1
FRT_TRACE_SORT = 2
TRACE_NOALLOC_MAX_RAYS = 1 << 20  # FRT_TRACE_NOALLOC_MAX_RAYS

# every exported symbol of include/frt.h and include/frt_host.h (tests check the libraries export them)
FRT_SYMBOLS = [
    "frt_create", "frt_destroy", "frt_last_error", "frt_set_flags", "frt_device_count",
    "frt_upload_scene", "frt_render_tile", "frt_clear_film", "frt_readback", "frt_film_device_ptr",
    "frt_trace_rays", "frt_trace_rays_device", "frt_stream_handle", "frt_get_counters", "frt_restore_film",
    "frt_readback_film", "frt_cancel", "frt_poll", "frt_host_alloc", "frt_host_free", "frt_build_wide8",
    "frt_trace_rays_device_noalloc",
]
FRTH_SYMBOLS = [
    "frth_last_error", "frth_scene_create", "frth_scene_create_soup", "frth_scene_destroy",
    "frth_scene_desc", "frth_scene_info", "frth_soup_triangles", "frth_camera_rays", "frth_render",
    "frth_bvh4_build", "frth_cancel", "frth_last_ldr", "frth_write_png",
]


class FrtError(RuntimeError):
    pass


class frt_tile(C.Structure):
    _fields_ = [(n, C.c_uint32) for n in
                ("nx", "ny", "x0", "y0", "w", "h", "spp_begin", "spp_end", "seed", "max_depth", "integrator", "reserved")]


FRT_INTEGRATOR_PATH, FRT_INTEGRATOR_AO, FRT_INTEGRATOR_NORMALS, FRT_INTEGRATOR_DIRECT = 0, 1, 2, 3
INTEGRATORS = {"path": FRT_INTEGRATOR_PATH, "ao": FRT_INTEGRATOR_AO, "normals": FRT_INTEGRATOR_NORMALS,
               "direct": FRT_INTEGRATOR_DIRECT}


class frt_stats(C.Structure):
    _fields_ = [("samples", C.c_uint64), ("closest_rays", C.c_uint64), ("shadow_rays", C.c_uint64),
                ("node_visits", C.c_uint64), ("tri_tests", C.c_uint64), ("iterations", C.c_uint64),
                ("kernel_launches", C.c_uint64), ("render_ms", C.c_double), ("extend_ms", C.c_double),
                ("shade_ms", C.c_double), ("connect_ms", C.c_double), ("generate_ms", C.c_double)]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


class frth_info(C.Structure):
    _fields_ = [("n_prims", C.c_uint32), ("n_tris", C.c_uint32), ("n_spheres", C.c_uint32),
                ("n_lights", C.c_uint32), ("n_materials", C.c_uint32), ("n_wide_nodes", C.c_uint32),
                ("wide_depth", C.c_uint32), ("bvh_nodes", C.c_uint32), ("bvh_leaves", C.c_uint32),
                ("build_ms", C.c_double), ("flatten_ms", C.c_double),
                ("bounds_lo", C.c_float * 3), ("bounds_hi", C.c_float * 3)]

    def as_dict(self):
        d = {n: getattr(self, n) for n, _ in self._fields_[:11]}
        d["bounds_lo"] = list(self.bounds_lo)
        d["bounds_hi"] = list(self.bounds_hi)
        return d


class frt_camera(C.Structure):
    _fields_ = [(n, C.c_float * 3) for n in
                ("origin", "lower_left_corner", "horizontal", "vertical", "u", "v", "w")] + [("lens_radius", C.c_float)]


class frt_scene_desc(C.Structure):
    _fields_ = [("struct_size", C.c_uint32), ("n_nodes", C.c_uint32), ("nodes", C.c_void_p),
                ("n_tris", C.c_uint32), ("n_spheres", C.c_uint32), ("tris", C.c_void_p),
                ("tri_normals", C.c_void_p), ("spheres", C.c_void_p), ("n_materials", C.c_uint32),
                ("n_lights", C.c_uint32), ("materials", C.c_void_p), ("lights", C.c_void_p),
                ("camera", frt_camera), ("env_color", C.c_float * 3), ("ao_t_max", C.c_float),
                ("env_image", C.c_void_p), ("env_nx", C.c_uint32), ("env_ny", C.c_uint32),
                ("n_textures", C.c_uint32), ("reserved0", C.c_uint32), ("textures", C.c_void_p),
                ("tex_texels", C.c_void_p), ("n_texels", C.c_uint64), ("tri_uv", C.c_void_p)]


RAY_DTYPE = np.dtype([("o", "<f4", 3), ("d", "<f4", 3)])
HIT_DTYPE = np.dtype([("t", "<f4"), ("b1", "<f4"), ("b2", "<f4"), ("prim_id", "<i4")])
NODE_DTYPE = np.dtype([("lo_x", "<f4", 4), ("hi_x", "<f4", 4), ("lo_y", "<f4", 4), ("hi_y", "<f4", 4),
                       ("lo_z", "<f4", 4), ("hi_z", "<f4", 4), ("child", "<u4", 4), ("reserved", "<u4", 4)])
MATERIAL_DTYPE = np.dtype([("type", "<u4"), ("p", "<f4", 15)])
LIGHT_DTYPE = np.dtype([("kind", "<u4"), ("material", "<u4"), ("flags", "<u4"), ("pad", "<u4"),
                        ("v0", "<f4", 4), ("v1", "<f4", 4), ("v2", "<f4", 4), ("n0", "<f4", 4),
                        ("n1", "<f4", 4), ("n2", "<f4", 4), ("reserved", "<f4", 4)])
TEXTURE_DTYPE = np.dtype([("kind", "<u4"), ("a", "<u4"), ("b", "<u4"), ("c", "<u4"), ("p", "<f4", 4)])
SPHERE_DTYPE = np.dtype([("c", "<f4", 3), ("radius", "<f4"), ("material", "<u4"), ("prim_id", "<u4"),
                         ("pad", "<u4", 2)])

_libs = {}


def lib_paths():
    return PKG_DIR / "libfrt.so", PKG_DIR / "libfrt_host.so"


def load_libs():
    """Load libfrt.so and libfrt_host.so from the package directory (built in-tree by build())."""
    if _libs:
        return _libs["frt"], _libs["host"]
    core, host = lib_paths()
    for p in (core, host):
        if not p.exists():
            raise FrtError(f"{p} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                           "(there is no Python/CPU fallback)")
    frt = C.CDLL(str(core), mode=C.RTLD_GLOBAL)
    frth = C.CDLL(str(host), mode=C.RTLD_GLOBAL)
    vp, u32, u64, sz, i32, dbl = C.c_void_p, C.c_uint32, C.c_uint64, C.c_size_t, C.c_int, C.c_double
    sig = {
        frt: {
            "frt_create": (i32, [C.POINTER(vp), C.POINTER(i32), i32]),
            "frt_destroy": (None, [vp]),
            "frt_last_error": (C.c_char_p, [vp]),
            "frt_set_flags": (i32, [vp, u32]),
            "frt_device_count": (i32, [vp]),
            "frt_upload_scene": (i32, [vp, vp]),
            "frt_render_tile": (i32, [vp, C.POINTER(frt_tile)]),
            "frt_clear_film": (i32, [vp]),
            "frt_readback": (i32, [vp, vp, sz, C.POINTER(frt_stats)]),
            "frt_get_counters": (i32, [vp, C.POINTER(frt_stats)]),
            "frt_restore_film": (i32, [vp, u32, u32, vp, sz]),
            "frt_film_device_ptr": (i32, [vp, i32, C.POINTER(vp), C.POINTER(sz)]),
            "frt_stream_handle": (i32, [vp, i32, C.POINTER(vp)]),
            "frt_trace_rays": (i32, [vp, vp, sz, vp, i32, C.POINTER(dbl)]),
            "frt_trace_rays_device": (i32, [vp, i32, vp, sz, vp, i32, C.POINTER(dbl)]),
            "frt_trace_rays_device_noalloc": (i32, [vp, i32, vp, sz, vp, i32, C.POINTER(dbl)]),
            "frt_readback_film": (i32, [vp, u64, vp, vp, vp, C.POINTER(frt_stats)]),
            "frt_cancel": (i32, [vp]),
            "frt_poll": (i32, [vp, C.POINTER(i32)]),
            "frt_host_alloc": (vp, [sz]),
            "frt_host_free": (None, [vp]),
            "frt_build_wide8": (i32, [vp, vp, sz, C.POINTER(u32), C.POINTER(i32)]),
        },
        frth: {
            "frth_last_error": (C.c_char_p, []),
            "frth_scene_create": (i32, [C.c_char_p, dbl, i32, C.POINTER(vp)]),
            "frth_scene_create_soup": (i32, [vp, sz, dbl, i32, C.POINTER(vp)]),
            "frth_scene_destroy": (None, [vp]),
            "frth_scene_desc": (C.POINTER(frt_scene_desc), [vp]),
            "frth_scene_info": (i32, [vp, C.POINTER(frth_info)]),
            "frth_soup_triangles": (i32, [sz, u64, C.c_float, vp]),
            "frth_camera_rays": (i32, [vp, i32, i32, vp]),
            "frth_render": (i32, [vp, i32, i32, i32, u32, u32, u32, vp, C.POINTER(frt_stats), C.POINTER(dbl)]),
            "frth_bvh4_build": (i32, [vp, sz, i32, dbl, C.POINTER(vp)]),
            "frth_cancel": (i32, []),
            "frth_last_ldr": (i32, [vp, sz]),
            "frth_write_png": (i32, [C.c_char_p, vp, i32, i32]),
        },
    }
    for lib, table in sig.items():
        for name, (res, args) in table.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
    _libs["frt"], _libs["host"] = frt, frth
    return frt, frth


def _ptr(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


class HostScene:
    """A scene built by the C++ host layer (reference-shaped classes + host SAH BVH) and flattened."""

    def __init__(self, name: str | None = None, aspect: float = 1.0, max_leaf: int = 4,
                 soup: np.ndarray | None = None, fast_bvh: bool = False):
        _, frth = load_libs()
        self._frth = frth
        self.handle = C.c_void_p()
        if "FRT_ASSETS" not in os.environ:
            os.environ["FRT_ASSETS"] = str(REPO_DIR / "assets")
        if soup is not None:
            soup = np.ascontiguousarray(soup, dtype=np.float32).reshape(-1, 9)
            if fast_bvh:
                rc = frth.frth_bvh4_build(_ptr(soup), soup.shape[0], max_leaf, aspect, C.byref(self.handle))
            else:
                rc = frth.frth_scene_create_soup(_ptr(soup), soup.shape[0], aspect, max_leaf, C.byref(self.handle))
        else:
            rc = frth.frth_scene_create(name.encode(), aspect, max_leaf, C.byref(self.handle))
        if rc != FRT_OK:
            raise FrtError(f"host scene '{name}': {frth.frth_last_error().decode()}")
        self.name = name or "soup"
        self.aspect = aspect

    def close(self):
        if self.handle:
            self._frth.frth_scene_destroy(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def desc(self) -> frt_scene_desc:
        return self._frth.frth_scene_desc(self.handle).contents

    @property
    def info(self) -> dict:
        i = frth_info()
        self._frth.frth_scene_info(self.handle, C.byref(i))
        return i.as_dict()

    def _view(self, ptr, count, dtype):
        if not ptr or count == 0:
            return np.zeros(0, dtype=dtype)
        buf = (C.c_char * (count * np.dtype(dtype).itemsize)).from_address(ptr)
        return np.frombuffer(buf, dtype=dtype).copy()

    def arrays(self) -> dict:
        """Copies of the flattened POD arrays (for host-side tests)."""
        d = self.desc
        return {
            "nodes": self._view(d.nodes, d.n_nodes, NODE_DTYPE),
            "tris": self._view(d.tris, d.n_tris * 3 * 4, np.float32).reshape(-1, 3, 4),
            "tri_normals": self._view(d.tri_normals, d.n_tris * 3 * 4, np.float32).reshape(-1, 3, 4),
            "spheres": self._view(d.spheres, d.n_spheres, SPHERE_DTYPE),
            "materials": self._view(d.materials, d.n_materials, MATERIAL_DTYPE),
            "lights": self._view(d.lights, d.n_lights, LIGHT_DTYPE),
            "textures": self._view(d.textures, d.n_textures, TEXTURE_DTYPE),
            "tex_texels": self._view(d.tex_texels, d.n_texels * 3, np.float32).reshape(-1, 3),
            "tri_uv": self._view(d.tri_uv, d.n_tris * 6 if d.tri_uv else 0, np.float32).reshape(-1, 3, 2),
        }

    def camera_rays(self, nx: int, ny: int) -> np.ndarray:
        rays = np.zeros(nx * ny, dtype=RAY_DTYPE)
        rc = self._frth.frth_camera_rays(self.handle, nx, ny, _ptr(rays))
        if rc != FRT_OK:
            raise FrtError("frth_camera_rays failed")
        return rays

    def render(self, nx: int, ny: int, spp: int, seed: int = 1, n_gpus: int = 1, flags: int = 0, integrator: str = "path"):
        """renderer<gpu_{path,ao,normals}_integrator>::Render → (mean image [ny,nx,3] row 0 = bottom, stats, seconds)."""
        flags = (flags & 0xFF) | (INTEGRATORS[integrator] << 8)
        img = np.zeros((ny, nx, 3), dtype=np.float32)
        st = frt_stats()
        secs = C.c_double()
        rc = self._frth.frth_render(self.handle, n_gpus, nx, ny, spp, seed, flags, _ptr(img), C.byref(st), C.byref(secs))
        if rc not in (FRT_OK, FRTH_CANCELLED):
            raise FrtError(f"frth_render: {self._frth.frth_last_error().decode()}")
        d = st.as_dict()
        d["cancelled"] = rc == FRTH_CANCELLED
        return img, d, secs.value

    def last_ldr(self, nx: int, ny: int) -> np.ndarray:
        """viewer::out_image of the last render: the 8-bit tone-mapped film [ny, nx, 3], row 0 = bottom."""
        out = np.zeros((ny, nx, 3), np.uint8)
        if self._frth.frth_last_ldr(_ptr(out), out.size) != FRT_OK:
            raise FrtError("frth_last_ldr: no film of that size")
        return out


def write_png(path: str, ldr: np.ndarray):
    """frth_write_png: an 8-bit RGB film [ny, nx, 3] with row 0 = bottom (HostScene.last_ldr) as a PNG file."""
    _, frth = load_libs()
    ldr = np.ascontiguousarray(ldr, dtype=np.uint8)
    assert ldr.ndim == 3 and ldr.shape[2] == 3
    if frth.frth_write_png(str(path).encode(), _ptr(ldr), ldr.shape[1], ldr.shape[0]) != FRT_OK:
        raise FrtError(f"frth_write_png: {frth.frth_last_error().decode()}")


def soup_triangles(n: int, seed: int = 0x5EED0001, edge: float = 0.004) -> np.ndarray:
    _, frth = load_libs()
    out = np.zeros((n, 9), dtype=np.float32)
    frth.frth_soup_triangles(n, seed, edge, _ptr(out))
    return out


class Context:
    """frt_ctx: one or more CUDA devices with an uploaded scene."""

    def __init__(self, devices=None, n: int = 1):
        frt, _ = load_libs()
        self._frt = frt
        self.handle = C.c_void_p()
        if devices is not None:
            arr = (C.c_int * len(devices))(*devices)
            rc = frt.frt_create(C.byref(self.handle), arr, len(devices))
        else:
            rc = frt.frt_create(C.byref(self.handle), None, n)
        if rc != FRT_OK:
            raise FrtError(f"frt_create: {frt.frt_last_error(None).decode()}")
        self.nx = self.ny = 0

    def _check(self, rc, what):
        if rc != FRT_OK:
            raise FrtError(f"{what}: {self._frt.frt_last_error(self.handle).decode()} (status {rc})")

    def close(self):
        if self.handle:
            self._frt.frt_destroy(self.handle)
            self.handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_flags(self, flags: int):
        self._check(self._frt.frt_set_flags(self.handle, flags), "frt_set_flags")

    def upload(self, scene: HostScene):
        self._scene = scene  # keep host arrays alive during the call
        self._check(self._frt.frt_upload_scene(self.handle, C.addressof(scene.desc)), "frt_upload_scene")

    def upload_desc(self, desc: frt_scene_desc):
        self._check(self._frt.frt_upload_scene(self.handle, C.addressof(desc)), "frt_upload_scene")

    def render_tile(self, nx, ny, spp_begin, spp_end, seed=1, max_depth=34, rect=None, integrator="path"):
        x0, y0, w, h = rect if rect else (0, 0, nx, ny)
        t = frt_tile(nx, ny, x0, y0, w, h, spp_begin, spp_end, seed, max_depth, INTEGRATORS[integrator], 0)
        self._check(self._frt.frt_render_tile(self.handle, C.byref(t)), "frt_render_tile")
        self.nx, self.ny = nx, ny

    def clear_film(self):
        self._check(self._frt.frt_clear_film(self.handle), "frt_clear_film")

    def readback(self, out: np.ndarray | None = None, want_image: bool = True):
        """→ (radiance SUMS [ny,nx,3] float32 row 0 = bottom, stats dict)."""
        st = frt_stats()
        if want_image:
            if out is None:
                out = np.zeros((self.ny, self.nx, 3), dtype=np.float32)
            self._check(self._frt.frt_readback(self.handle, _ptr(out), out.size, C.byref(st)), "frt_readback")
        else:
            self._check(self._frt.frt_readback(self.handle, None, 0, C.byref(st)), "frt_readback")
        return out, st.as_dict()

    def restore_film(self, sums: np.ndarray):
        """Film sums [ny,nx,3] float32 from an earlier readback become the film's starting state (resume)."""
        sums = np.ascontiguousarray(sums, dtype=np.float32)
        ny, nx = sums.shape[:2]
        self._check(self._frt.frt_restore_film(self.handle, nx, ny, _ptr(sums), sums.size), "frt_restore_film")
        self.nx, self.ny = nx, ny

    def counters(self) -> dict:
        st = frt_stats()
        self._check(self._frt.frt_get_counters(self.handle, C.byref(st)), "frt_get_counters")
        return st.as_dict()

    def film_device_ptr(self, dev: int = 0):
        p, n = C.c_void_p(), C.c_size_t()
        self._check(self._frt.frt_film_device_ptr(self.handle, dev, C.byref(p), C.byref(n)), "frt_film_device_ptr")
        return p.value, n.value

    def stream_handle(self, dev: int = 0) -> int:
        p = C.c_void_p()
        self._check(self._frt.frt_stream_handle(self.handle, dev, C.byref(p)), "frt_stream_handle")
        return p.value

    def trace(self, rays: np.ndarray, any_hit: bool = False, sort: bool = False, hits: np.ndarray | None = None):
        """Host buffers in, host buffers out (H2D / D2H inside the call) → (hits, device_ms)."""
        rays = np.ascontiguousarray(rays)
        assert rays.dtype == RAY_DTYPE
        if hits is None:
            hits = np.zeros(rays.shape[0], dtype=HIT_DTYPE)
        ms = C.c_double()
        mode = (FRT_TRACE_ANY_HIT if any_hit else 0) | (FRT_TRACE_SORT if sort else 0)
        self._check(self._frt.frt_trace_rays(self.handle, _ptr(rays), rays.shape[0], _ptr(hits), mode, C.byref(ms)), "frt_trace_rays")
        return hits, ms.value

    def trace_device(self, d_rays_ptr: int, n: int, d_hits_ptr: int, any_hit: bool = False, dev: int = 0, sort: bool = False) -> float:
        ms = C.c_double()
        mode = (FRT_TRACE_ANY_HIT if any_hit else 0) | (FRT_TRACE_SORT if sort else 0)
        self._check(self._frt.frt_trace_rays_device(self.handle, dev, d_rays_ptr, n, d_hits_ptr, mode, C.byref(ms)), "frt_trace_rays_device")
        return ms.value

    def trace_device_noalloc(self, d_rays_ptr: int, n: int, d_hits_ptr: int, any_hit: bool = False, dev: int = 0) -> float:
        """frt_trace_rays_device_noalloc: the Morton-ordered cast over the library's static scratch (n <= TRACE_NOALLOC_MAX_RAYS)."""
        ms = C.c_double()
        self._check(self._frt.frt_trace_rays_device_noalloc(self.handle, dev, d_rays_ptr, n, d_hits_ptr,
                                                            FRT_TRACE_ANY_HIT if any_hit else 0, C.byref(ms)), "frt_trace_rays_device_noalloc")
        return ms.value

    def readback_film(self, ns: int):
        """frt_readback_film → (mean [ny,nx,3] float64 = viewer::fout_image, 8-bit tone-mapped image = viewer::out_image, stats)."""
        st = frt_stats()
        fout = np.zeros((self.ny, self.nx, 3), dtype=np.float64)
        ldr = np.zeros((self.ny, self.nx, 3), dtype=np.uint8)
        self._check(self._frt.frt_readback_film(self.handle, ns, _ptr(fout), _ptr(ldr), None, C.byref(st)), "frt_readback_film")
        return fout, ldr, st.as_dict()

    def cancel(self):
        self._check(self._frt.frt_cancel(self.handle), "frt_cancel")

    def poll(self) -> bool:
        done = C.c_int()
        self._check(self._frt.frt_poll(self.handle, C.byref(done)), "frt_poll")
        return bool(done.value)


class Pinned:
    """A numpy array over frt_host_alloc'ed (pinned) host memory: `.array`; `.close()` frees it (drop every view first)."""

    def __init__(self, n: int, dtype):
        frt, _ = load_libs()
        self._frt = frt
        dt = np.dtype(dtype)
        self.nbytes = max(1, n * dt.itemsize)
        self.ptr = frt.frt_host_alloc(self.nbytes)
        if not self.ptr:
            raise FrtError("frt_host_alloc failed")
        self.array = np.frombuffer((C.c_char * self.nbytes).from_address(self.ptr), dtype=dt, count=n)

    def close(self):
        if self.ptr:
            self.array = None
            self._frt.frt_host_free(self.ptr)
            self.ptr = None


def build_wide8(scene: "HostScene"):
    """frt_build_wide8 → (words [n8, 24] uint32, depth)."""
    frt, _ = load_libs()
    n8, depth = C.c_uint32(), C.c_int()
    rc = frt.frt_build_wide8(C.addressof(scene.desc), None, 0, C.byref(n8), C.byref(depth))
    if rc != FRT_OK:
        raise FrtError("frt_build_wide8 failed")
    words = np.zeros((n8.value, 24), np.uint32)
    rc = frt.frt_build_wide8(C.addressof(scene.desc), _ptr(words), words.size, C.byref(n8), C.byref(depth))
    if rc != FRT_OK:
        raise FrtError("frt_build_wide8 failed")
    return words, depth.value


def save_checkpoint(path, sums: np.ndarray, spp_done: int, seed: int, scene: str = "", integrator: str = "path"):
    """Film checkpoint of a long render: the radiance sums and how many samples per pixel they hold.  Because a
    sample is identified by (seed, pixel, index), rendering [spp_done, spp) on top of the restored sums gives the
    film an uninterrupted render would have given, on any number of GPUs."""
    np.savez_compressed(path, sums=np.asarray(sums, dtype=np.float32), spp_done=np.int64(spp_done), seed=np.int64(seed),
                        scene=np.str_(scene), integrator=np.str_(integrator))


def load_checkpoint(path) -> dict:
    with np.load(path) as z:
        return {"sums": z["sums"], "spp_done": int(z["spp_done"]), "seed": int(z["seed"]), "scene": str(z["scene"]),
                "integrator": str(z["integrator"])}
