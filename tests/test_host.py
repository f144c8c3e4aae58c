"""CPU tests of the host layer and of the C-ABI surface (no compute calls: there is no GPU here).

 * both shared libraries load and export every symbol include/*.h declares;
 * without a CUDA device the product fails loudly (no CPU fallback);
 * the flattened scene description equals what the reference driver builds for the same scene;
 * the wide BVH is well formed: every primitive in exactly one leaf, child boxes enclose them;
 * the fast soup builder and the reference-shaped builder describe the same triangles."""
import ctypes as C
import re

import numpy as np
import pytest

import refutil
from first_raytracer_b200 import api

REPO = refutil.REPO


def _declared(header):
    text = (REPO / "include" / header).read_text()
    return sorted(set(re.findall(r"FRT_API\s+[\w\s\*]+?\b(frth?_\w+)\s*\(", text)))


def test_abi_exports_every_declared_symbol(built):
    frt, frth = api.load_libs()
    core, host = _declared("frt.h"), _declared("frt_host.h")
    assert len(core) >= 12 and len(host) >= 9
    for name in core:
        assert hasattr(frt, name), f"libfrt.so does not export {name}"
    for name in host:
        assert hasattr(frth, name), f"libfrt_host.so does not export {name}"
    assert sorted(api.FRT_SYMBOLS) == core
    assert sorted(api.FRTH_SYMBOLS) == host


def test_struct_layouts_match_header(built):
    assert C.sizeof(api.frt_tile) == 48
    assert api.NODE_DTYPE.itemsize == 128 and api.MATERIAL_DTYPE.itemsize == 64
    assert api.LIGHT_DTYPE.itemsize == 128 and api.SPHERE_DTYPE.itemsize == 32
    assert api.RAY_DTYPE.itemsize == 24 and api.HIT_DTYPE.itemsize == 16
    scene = api.HostScene("cornell", 1.0, 4)
    assert scene.desc.struct_size == C.sizeof(api.frt_scene_desc)


def test_no_cpu_fallback(built):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(api.FrtError, match="no CUDA device|CUDA"):
        api.Context()
    scene = api.HostScene("cornell", 1.0, 4)
    with pytest.raises(api.FrtError):
        scene.render(8, 8, 1)


def _leaf_ranges(nodes):
    tri_seen, sph_seen = [], []
    for n in nodes:
        for c in n["child"]:
            c = int(c)
            if c == 0xFFFFFFFF or not (c & 0x80000000):
                continue
            first, count = c & 0x03FFFFFF, ((c >> 26) & 15) + 1
            (sph_seen if c & 0x40000000 else tri_seen).extend(range(first, first + count))
    return tri_seen, sph_seen


@pytest.mark.parametrize("name,aspect,max_leaf", [("cornell", 1.0, 1), ("cornell", 1.0, 4), ("veach", 1.5, 4),
                                                  ("teapot", 16 / 9, 4), ("teapot", 16 / 9, 8), ("cornell_sphere", 1.0, 2)])
def test_wide_bvh_is_well_formed(built, name, aspect, max_leaf):
    scene = api.HostScene(name, aspect, max_leaf)
    a = scene.arrays()
    nodes, tris = a["nodes"], a["tris"]
    tri_seen, sph_seen = _leaf_ranges(nodes)
    assert sorted(tri_seen) == list(range(tris.shape[0])), "every triangle in exactly one leaf"
    assert sorted(sph_seen) == list(range(a["spheres"].shape[0]))
    ids = np.concatenate([tris[:, 0, 3].view(np.uint32), a["spheres"]["prim_id"]])
    assert sorted(ids.tolist()) == list(range(ids.size)), "creation-order ids are a permutation"
    # children follow parents (frt_upload_scene's validation relies on it) and boxes enclose the geometry
    for i, n in enumerate(nodes):
        for k in range(4):
            c = int(n["child"][k])
            if c == 0xFFFFFFFF:
                continue
            lo = np.array([n["lo_x"][k], n["lo_y"][k], n["lo_z"][k]])
            hi = np.array([n["hi_x"][k], n["hi_y"][k], n["hi_z"][k]])
            if not (c & 0x80000000):
                assert i < c < len(nodes)
                ch = nodes[c]
                used = ch["child"] != 0xFFFFFFFF
                clo = np.array([ch["lo_x"][used].min(), ch["lo_y"][used].min(), ch["lo_z"][used].min()])
                chi = np.array([ch["hi_x"][used].max(), ch["hi_y"][used].max(), ch["hi_z"][used].max()])
                assert (clo >= lo - 1e-4 * np.abs(lo) - 1e-5).all() and (chi <= hi + 1e-4 * np.abs(hi) + 1e-5).all()
            elif not (c & 0x40000000):
                first, count = c & 0x03FFFFFF, ((c >> 26) & 15) + 1
                assert count <= max_leaf
                t = tris[first:first + count]
                v0, v1, v2 = t[:, 0, :3], t[:, 1, :3], t[:, 2, :3]
                pts = np.concatenate([v0, v1, v2]).astype(np.float64)
                tol = 1e-5 * max(1.0, np.abs(pts).max())
                assert (pts >= lo - tol).all() and (pts <= hi + tol).all(), "leaf box must enclose its triangles"


@pytest.mark.skipif(not refutil.have_ref(), reason="oracle/_ref not built (needs /root/reference)")
@pytest.mark.parametrize("name,aspect", [("cornell", 1.0), ("veach", 1.5), ("teapot", 16 / 9), ("cornell_sphere", 1.0),
                                         ("cornell_glossy", 1.0)])
def test_scene_description_equals_reference_dump(built, name, aspect):
    """The product's own OBJ/MTL loader + scene builders feed the GPU what the reference's own loader produces:
    oracle/_ref/frt_ref reads the same OBJ files through the reference's mesh_loader.cpp and its vendored Assimp 5.0.1
    (number parsing, quad splitting, GenSmoothNormals, material names), so this pins host/mesh_loader.cpp on it."""
    ref, info = refutil.ref_dump(name, aspect)
    scene = api.HostScene(name, aspect, 4)
    a = scene.arrays()
    tris = a["tris"]
    order = np.argsort(tris[:, 0, 3].view(np.uint32))
    is_tri = ref["kind"] == 0
    g = ref["geom"][is_tri]
    for k in range(3):  # float32 vertices are exact (the description stores v0, v1, v2)
        assert np.abs(tris[order, k, :3] - g[:, k]).max() == 0.0
    if a["tri_normals"].size:
        assert np.abs(a["tri_normals"][order][:, :, :3] - g[:, 3:6]).max() < 1e-6
    mats = a["materials"]
    mid = tris[order, 1, 3].view(np.uint32)
    assert (mats["type"][mid] == ref["mat_kind"][is_tri]).all()
    assert np.abs(mats["p"][mid][:, :11] - ref["mat_p"][is_tri][:, :11]).max() < 1e-5
    geo_flag = (tris[order, 2, 3].view(np.uint32) & 1).astype(np.float64)
    assert (geo_flag == ref["geo_normals"][is_tri]).all()
    assert scene.info["n_lights"] == ref["n_lights"]
    cam = scene.desc.camera
    mine = np.concatenate([np.array(getattr(cam, f)) for f in
                           ("origin", "lower_left_corner", "horizontal", "vertical", "u", "v", "w")] + [[cam.lens_radius]])
    assert np.abs(mine - ref["camera"]).max() < 1e-5 * max(1.0, np.abs(ref["camera"]).max())
    # the host SAH build reports the node counts the reference prints (Scene::num_bvh_nodes, parallel_bvh.h:70,132,146)
    assert scene.info["bvh_nodes"] == info["bvh_nodes"] and scene.info["bvh_leaves"] == info["bvh_leaves"]


def test_soup_generator_and_builders(built):
    tri = api.soup_triangles(5000)
    assert np.array_equal(tri, api.soup_triangles(5000)), "splitmix64 stream is deterministic"
    assert tri.min() >= -0.004 and tri.max() <= 1.004
    slow = api.HostScene(soup=tri, max_leaf=4)
    fast = api.HostScene(soup=tri, max_leaf=4, fast_bvh=True)
    for s in (slow, fast):
        a = s.arrays()
        ids = a["tris"][:, 0, 3].view(np.uint32)
        order = np.argsort(ids)
        assert np.array_equal(np.sort(ids), np.arange(5000))
        assert np.array_equal(a["tris"][order][:, :, :3].reshape(-1, 9), tri)
        seen, _ = _leaf_ranges(a["nodes"])
        assert sorted(seen) == list(range(5000))


def test_invalid_arguments_are_reported(built):
    with pytest.raises(api.FrtError, match="unknown scene"):
        api.HostScene("no_such_scene", 1.0, 4)


def test_checkpoint_file_round_trip(tmp_path):
    """api.save_checkpoint / load_checkpoint keep the sums bit for bit and the bookkeeping that makes a film resumable."""
    rng = np.random.default_rng(3)
    sums = rng.random((6, 9, 3), dtype=np.float32) * 100
    api.save_checkpoint(tmp_path / "f.npz", sums, spp_done=17, seed=4, scene="veach", integrator="ao")
    ck = api.load_checkpoint(tmp_path / "f.npz")
    assert ck["sums"].dtype == np.float32 and np.array_equal(ck["sums"], sums)
    assert (ck["spp_done"], ck["seed"], ck["scene"], ck["integrator"]) == (17, 4, "veach", "ao")


def _write_hdr(path, rgbe, rle):
    """A Radiance picture from [ny, nx, 4] uint8 RGBE texels: flat, or the run-length coded scanlines of
    FORMAT=32-bit_rle_rgbe (per channel: count > 128 = a run of count - 128 equal bytes, else count literal bytes)."""
    ny, nx, _ = rgbe.shape
    out = bytearray(b"#?RADIANCE\n# test\nFORMAT=32-bit_rle_rgbe\n\n" + f"-Y {ny} +X {nx}\n".encode())
    if not rle:
        out += rgbe.tobytes()
    else:
        for j in range(ny):
            out += bytes([2, 2, nx >> 8, nx & 255])
            for k in range(4):
                row, i = rgbe[j, :, k], 0
                while i < nx:
                    run = 1
                    while i + run < nx and run < 127 and row[i + run] == row[i]:
                        run += 1
                    if run >= 3:
                        out += bytes([128 + run, int(row[i])])
                        i += run
                    else:
                        lit = min(nx - i, 5)
                        out += bytes([lit]) + row[i:i + lit].tobytes()
                        i += lit
    path.write_bytes(bytes(out))


@pytest.mark.parametrize("nx,ny,rle", [(4, 3, False), (16, 5, True), (40, 6, True)])
def test_radiance_hdr_reader_and_environment_map(built, tmp_path, monkeypatch, nx, ny, rle):
    """environment_map(filename) (material.h:209-213) through the host layer's own Radiance reader: the floats handed
    to frt_scene_desc::env_image are mantissa * 2^(exponent - 136) as stbi_loadf (image.cpp:16) produces, exponent 0 =
    black, row 0 = top — for flat and for run-length coded files (the reference's stb-loaded data/test.hdr is
    covered by the +hdr golden renders)."""
    rng = np.random.default_rng(nx * 100 + ny)
    rgbe = rng.integers(0, 256, (ny, nx, 4), dtype=np.uint8)
    rgbe[:, :, 3] = rng.integers(120, 140, (ny, nx))
    rgbe[0, : nx // 2] = rgbe[0, 0]  # a long run
    rgbe[1, 1, 3] = 0                # exponent 0: black whatever the mantissas
    for sub in ("CornellBox", "cube", "veach_mi"):
        (tmp_path / sub).symlink_to(REPO / "assets" / sub)
    (tmp_path / "data").mkdir()
    _write_hdr(tmp_path / "data" / "test.hdr", rgbe, rle)
    monkeypatch.setenv("FRT_ASSETS", str(tmp_path))
    scene = api.HostScene("cornell+hdr", 1.0, 2)
    d = scene.desc
    assert (d.env_nx, d.env_ny) == (nx, ny) and d.env_image
    got = np.ctypeslib.as_array(C.cast(d.env_image, C.POINTER(C.c_float)), shape=(ny, nx, 3)).copy()
    want = rgbe[:, :, :3].astype(np.float32) * np.ldexp(np.float32(1), rgbe[:, :, 3:].astype(np.int32) - 136).astype(np.float32)
    want[rgbe[:, :, 3] == 0] = 0
    assert np.array_equal(got, want)
    scene.close()
    # a truncated file is an error, not a partly filled image
    full = (tmp_path / "data" / "test.hdr").read_bytes()
    (tmp_path / "data" / "test.hdr").write_bytes(full[:-7])
    with pytest.raises(api.FrtError):
        api.HostScene("cornell+hdr", 1.0, 2)


def test_constant_environment_has_no_image(built):
    scene = api.HostScene("cornell+sky", 1.0, 2)
    assert not scene.desc.env_image and scene.desc.env_nx == 0
    assert np.allclose(list(scene.desc.env_color), [0.5, 0.6, 0.8])


def test_textured_scene_description(built):
    """Surface textures (texture.h:30-112) reach the C ABI as a table: checker children and image extents stay inside
    their arrays, the LDR texels are FromSrgb(byte / 255) of assets/data/tiles.bmp in the reference's row order
    (stb_image with vertical flip, image.cpp:19-21 = the BMP file's own bottom-up order), triangles carry mesh uv."""
    scene = api.HostScene("textured", 1.5, 2)
    a = scene.arrays()
    tex, texels = a["textures"], a["tex_texels"]
    assert scene.desc.struct_size == C.sizeof(api.frt_scene_desc)
    kinds = set(tex["kind"].tolist())
    assert kinds == {0, 1, 2}
    for t in tex:
        if t["kind"] == 1:
            assert t["a"] < len(tex) and t["b"] < len(tex)
        if t["kind"] == 2:
            assert (t["b"], t["c"]) == (64, 32) and t["a"] + 64 * 32 <= len(texels)
    raw = (REPO / "assets" / "data" / "tiles.bmp").read_bytes()
    bgr = np.frombuffer(raw[54:], np.uint8).reshape(32, 64, 3).astype(np.float64) / 255.0
    rgb = bgr[:, :, ::-1]
    lin = np.where(rgb <= 0.04045, rgb / 12.92, ((rgb + 0.055) / 1.055) ** 2.4)
    first = int(tex[tex["kind"] == 2][0]["a"])
    assert np.allclose(texels[first:first + 64 * 32].reshape(32, 64, 3), lin, rtol=1e-6, atol=1e-7)
    # textured materials carry 1 + texture index in p[11] / p[12]; constants keep 0 there
    refs = a["materials"]["p"][:, 11:13].copy().view(np.uint32)
    assert refs.max() <= len(tex) and (refs > 0).sum() >= 5
    uv = a["tri_uv"]
    assert uv.shape == (scene.desc.n_tris, 3, 2) and uv.min() == 0.0 and uv.max() == 1.0
    # a scene of constant colours uploads no texture arrays at all
    plain = api.HostScene("cornell", 1.0, 2)
    assert plain.desc.n_textures == 0 and not plain.desc.tri_uv and not plain.desc.tex_texels


def test_round2_scene_variants_equal_reference_driver(built):
    """cornell_metal / cornell_lens / veach_beckmann / textured: the host layer's scenes hold what the reference driver
    (oracle/ref_driver/ref_main.cpp, the reference's own classes) builds under the same names."""
    if not refutil.have_ref():
        pytest.skip("oracle/_ref not built")
    for name, aspect in [("cornell_metal", 1.0), ("cornell_lens", 1.0), ("veach_beckmann", 1.5), ("textured", 1.5)]:
        scene = api.HostScene(name, aspect, 2)
        ref, info = refutil.ref_dump(name, aspect)
        a = scene.arrays()
        d = scene.desc
        cam = np.concatenate([np.asarray(getattr(d.camera, f)[:], np.float64) for f in
                              ("origin", "lower_left_corner", "horizontal", "vertical", "u", "v", "w")] + [[d.camera.lens_radius]])
        assert np.allclose(cam, ref["camera"], rtol=1e-6, atol=1e-6), name
        assert d.n_tris + d.n_spheres == len(ref["kind"]) and d.n_lights == ref["n_lights"]
        # material classes per primitive (creation order): the reference's dynamic type vs the flattened tag
        ids = a["tris"][:, 0, 3].view(np.uint32)
        mats = a["tris"][:, 1, 3].view(np.uint32)
        for i, m in zip(ids, mats):
            assert int(ref["mat_kind"][i]) == int(a["materials"]["type"][m]), (name, i)
        for s in a["spheres"]:
            assert int(ref["mat_kind"][s["prim_id"]]) == int(a["materials"]["type"][s["material"]]), name
        if name == "veach_beckmann":
            rc = a["materials"][a["materials"]["type"] == 3]
            assert len(rc) == 4 and (rc["p"][:, 10] == 1.0).all()
        if name == "cornell_lens":
            assert abs(d.camera.lens_radius - 0.1) < 1e-7


@pytest.mark.parametrize("name,aspect,max_leaf", [("cornell", 1.0, 2), ("veach", 1.5, 2), ("teapot", 16 / 9, 2), ("cornell_sphere", 1.0, 4),
                                                  ("cornell_glossy", 1.0, 2), ("cornell_metal", 1.0, 2),
                                                  ("textured", 1.5, 2), ("cornell_lens", 1.0, 2), ("veach_beckmann+sky", 1.5, 2)])
def test_drop_in_build_flattens_the_reference_scene_identically(built, name, aspect, max_leaf):
    """The drop-in claim of INTEGRATION.md, guarded: host/flatten.h + host/gpu_path_integrator.h compile with
    -DFRT_WITH_REFERENCE_HEADERS against the reference's own headers (oracle/Makefile target `dropin`), and the
    flattener, walking the REFERENCE's Scene (its parallel_bvh_node / triangle / sphere / material / texture objects,
    built by the reference's own create_bvh), produces byte for byte the description the host layer produces."""
    if not refutil.DROPIN_BIN.exists():
        pytest.skip("oracle/_ref/frt_dropin not built (needs /root/reference)")
    ref = refutil.dropin_flatten(name, aspect, max_leaf)
    scene = api.HostScene(name, aspect, max_leaf)
    d = scene.desc

    def raw(ptr, nbytes):
        return bytes((C.c_char * nbytes).from_address(ptr)) if ptr and nbytes else b""

    assert raw(d.nodes, 128 * d.n_nodes) == ref["nodes"], "wide nodes differ"
    assert raw(d.tris, 48 * d.n_tris) == ref["tris"], "triangles differ"
    if name == "teapot":
        # The one array that cannot be byte-equal: teapot.obj has no normals, so they are Assimp's GenSmoothNormals on
        # the reference side — a float sum over the corners at a position, taken in the order of an unstable std::sort
        # (GenVertexNormalsProcess.cpp:190-196) — and the host loader's restatement of it.  Last-bit differences only.
        mine = np.frombuffer(raw(d.tri_normals, 48 * d.n_tris), np.float32)
        theirs = np.frombuffer(ref["tri_normals"], np.float32)
        assert mine.shape == theirs.shape and np.abs(mine - theirs).max() <= 2.5e-7
    else:
        assert raw(d.tri_normals, 48 * d.n_tris) == ref["tri_normals"]
    assert raw(d.spheres, 32 * d.n_spheres) == ref["spheres"]
    assert raw(d.materials, 64 * d.n_materials) == ref["materials"]
    assert raw(d.lights, 128 * d.n_lights) == ref["lights"]
    assert bytes(d.camera) == ref["camera"]
    assert bytes(d.env_color) == ref["env_color"]
    assert raw(d.textures, 32 * d.n_textures) == ref["textures"]
    assert raw(d.tex_texels, 12 * d.n_texels) == ref["tex_texels"]
    assert raw(d.tri_uv, 24 * d.n_tris) == ref["tri_uv"]
    assert bool(d.env_image) == ref["has_env_image"]


def _decode_wide8(words):
    """csrc/frt_wide8.cuh layout -> per node: origin (3), spacing (3), lo/hi planes [3][8] in world units, child refs [8]."""
    w = np.ascontiguousarray(words)
    origin = w[:, 0:3].copy().view(np.float32)
    exps = w[:, 3]
    spacing = np.stack([np.ldexp(1.0, ((exps >> (8 * a)) & 0xFF).astype(np.int64) - 127) for a in range(3)], axis=1)
    planes = w[:, 4:16].copy().view(np.uint8).reshape(-1, 6, 8).astype(np.float64)
    lo = origin[:, :, None].astype(np.float64) + planes[:, 0:3] * spacing[:, :, None]
    hi = origin[:, :, None].astype(np.float64) + planes[:, 3:6] * spacing[:, :, None]
    return lo, hi, w[:, 16:24]


@pytest.mark.parametrize("name,aspect,soup", [("teapot", 16 / 9, 0), ("cornell_sphere", 1.0, 0), ("textured", 1.5, 0), (None, 1.0, 50_000)])
def test_wide8_nodes_contain_the_wide4_tree(built, name, aspect, soup):
    """frt_build_wide8 (what frt_upload_scene derives for large scenes): every primitive range of the 4-wide tree is
    reachable exactly once, every child box CONTAINS the 4-wide tree's float32 box of the same subtree (so a walk of
    either tree decides the same candidates), inner children are wider than the 4-wide tree's (<= 8, mean > 4)."""
    scene = api.HostScene(soup=api.soup_triangles(soup), fast_bvh=True, max_leaf=4) if soup else api.HostScene(name, aspect, 2)
    a = scene.arrays()
    n4 = a["nodes"]
    words, depth = api.build_wide8(scene)
    lo, hi, child = _decode_wide8(words)
    assert depth >= 1 and 7 * depth + 1 <= 96
    # 4-wide boxes of every leaf reference and of every inner node (union of its children)
    def box4(i, k):
        return (np.array([n4[i]["lo_x"][k], n4[i]["lo_y"][k], n4[i]["lo_z"][k]], np.float64),
                np.array([n4[i]["hi_x"][k], n4[i]["hi_y"][k], n4[i]["hi_z"][k]], np.float64))
    leaf_box = {}
    for i in range(len(n4)):
        for k in range(4):
            c = int(n4[i]["child"][k])
            if c != 0xFFFFFFFF and c & 0x80000000:
                leaf_box[c] = box4(i, k)
    seen = []
    n_children = []
    stack = [0]
    visited = set()
    while stack:
        i = stack.pop()
        assert i not in visited
        visited.add(i)
        kids = 0
        for k in range(8):
            c = int(child[i, k])
            if c == 0xFFFFFFFF:
                assert (lo[i, :, k] > hi[i, :, k]).all(), "an empty slot must hold inverted planes"
                continue
            kids += 1
            if c & 0x80000000:
                seen.append(c)
                blo, bhi = leaf_box[c]
                assert (lo[i, :, k] <= blo).all() and (hi[i, :, k] >= bhi).all(), "leaf box not contained"
            else:
                assert c > i and c < len(words)
                stack.append(c)
        n_children.append(kids)
    assert len(visited) == len(words)
    assert sorted(seen) == sorted(leaf_box.keys()), "every leaf range exactly once"
    assert max(n_children) <= 8 and (len(n4) < 100 or np.mean(n_children) > 4.5)
    print(f"{name or 'soup'}: {len(n4)} 4-wide nodes ({128 * len(n4)} B) -> {len(words)} 8-wide nodes ({96 * len(words)} B), "
          f"depth {depth}, mean children {np.mean(n_children):.2f}")


def test_png_writer_round_trip(built, tmp_path):
    """frth_write_png (the headless stand-in for the reference's GL window): a valid PNG — signature, chunk CRCs, zlib
    stream with its Adler-32 — that decodes to the film with the rows flipped (row 0 of the film is the bottom row)."""
    import struct
    import zlib
    rng = np.random.default_rng(3)
    for ny, nx in ((1, 1), (7, 5), (300, 401)):  # the last one spans several 65535-byte stored blocks
        film = rng.integers(0, 256, (ny, nx, 3), dtype=np.uint8)
        path = tmp_path / f"film_{nx}x{ny}.png"
        api.write_png(path, film)
        raw = path.read_bytes()
        assert raw[:8] == b"\x89PNG\r\n\x1a\n"
        pos, chunks = 8, []
        while pos < len(raw):
            n, tag = struct.unpack(">I4s", raw[pos:pos + 8])
            body = raw[pos + 8:pos + 8 + n]
            assert struct.unpack(">I", raw[pos + 8 + n:pos + 12 + n])[0] == zlib.crc32(tag + body)
            chunks.append((tag, body))
            pos += 12 + n
        assert [t for t, _ in chunks] == [b"IHDR", b"IDAT", b"IEND"]
        assert struct.unpack(">IIBBBBB", chunks[0][1]) == (nx, ny, 8, 2, 0, 0, 0)
        rows = np.frombuffer(zlib.decompress(chunks[1][1]), np.uint8).reshape(ny, 1 + 3 * nx)
        assert (rows[:, 0] == 0).all()
        assert np.array_equal(rows[:, 1:].reshape(ny, nx, 3), film[::-1])
