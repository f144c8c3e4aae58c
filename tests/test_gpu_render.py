"""GPU parity, part 2: the wavefront path integrator against the reference's own path::Li.

Two kinds of comparison:
 * same random numbers: oracle/_ref/frt_ref_ctr is the UNMODIFIED path.cpp / material.h / pdf.h
   compiled with a counter-based `sampler` (oracle/ref_driver/ref_prelude.h), so reference and GPU
   consume identical uniforms per (pixel, sample, vertex).  Images then differ only by float32 vs
   float64 rounding (rare path flips), which makes a low-spp render a sharp test of every weight.
 * independent random numbers: the as-shipped mt19937 build, statistical bars of BASELINE.json
   (relMSE <= 1e-3 after block averaging at reduced spp, mean radiance within 0.5 %).
"""
import ctypes as C

import numpy as np
import pytest

import refutil
from first_raytracer_b200 import api

pytestmark = pytest.mark.gpu


def _gpu_render(name, nx, ny, spp, seed=1, max_leaf=4):
    scene = api.HostScene(name, nx / ny, max_leaf)
    ctx = api.Context()
    ctx.upload(scene)
    ctx.render_tile(nx, ny, 0, spp, seed=seed)
    sums, st = ctx.readback()
    ctx.close()
    return sums / spp, st


def _whole_image_mean_bar(gpu, ref, close, label):
    """The mean over ALL pixels (not only the agreeing ones): within BASELINE.json's 0.5 %.  The pixels that disagree
    are listed first, so that a bias hiding in them (a path that flipped onto an 800-radiance light) is visible."""
    dm = abs(gpu.mean() - ref.mean()) / ref.mean()
    bad = np.argwhere(~close)
    worst = [(int(y), int(x), float(gpu[y, x].max()), float(ref[y, x].max())) for y, x in bad[:4]]
    print(f"{label}: whole-image mean diff {dm:.3e} ({(~close).sum()} pixels outside 0.2 %: {worst})")
    assert dm < 5e-3, f"{label}: whole-image mean differs by {dm:.3%}"
    return dm


@pytest.mark.parametrize("name,nx,ny,spp", [("cornell", 128, 128, 16), ("veach", 192, 128, 16),
                                            ("teapot", 160, 90, 8), ("cornell_glossy", 96, 96, 8),
                                            ("cornell_sphere", 96, 96, 8),
                                            # round 2: Beckmann plates (pdf.h:280-403), a metal box (material.h:111-130), the thin
                                            # lens (camera.h:30-35), textured surfaces (texture.h:30-112)
                                            ("veach_beckmann", 192, 128, 16), ("cornell_metal", 96, 96, 8),
                                            ("cornell_lens", 96, 96, 8), ("textured", 192, 128, 8), ("random_scene", 192, 128, 8)])
def test_same_random_numbers_as_reference(built, name, nx, ny, spp):
    if not refutil.have_ref():
        pytest.skip("oracle/_ref not built")
    gpu, st = _gpu_render(name, nx, ny, spp, seed=7)
    ref, info = refutil.ref_render(name, nx, ny, spp, ctr=True, seed=7)
    assert np.isfinite(gpu).all()
    mean_gpu, mean_ref = gpu.mean(), ref.mean()
    # per-pixel agreement: most pixels identical up to float rounding
    diff = np.abs(gpu - ref).max(axis=2)
    scale = np.maximum(ref.max(axis=2), 1e-3)
    close = (diff / scale) < 2e-3
    print(f"{name}: mean gpu {mean_gpu:.6f} ref {mean_ref:.6f}, pixels within 0.2%: {close.mean():.4f}, "
          f"relMSE {refutil.rel_mse(gpu, ref):.3e}, gpu rays {st['closest_rays']}+{st['shadow_rays']}, "
          f"ref rays {info['closest_rays']}+{info['shadow_rays']}")
    assert abs(gpu[close].mean() - ref[close].mean()) / ref[close].mean() < 1e-4
    assert close.mean() > 0.999
    _whole_image_mean_bar(gpu, ref, close, name)
    # closest-ray counts must agree closely: same paths, same lengths
    assert abs(st["closest_rays"] - info["closest_rays"]) / info["closest_rays"] < 2e-3


@pytest.mark.parametrize("name,nx,ny,spp", [("cornell", 128, 128, 16), ("veach", 192, 128, 16), ("cornell", 48, 48, 4),
                                            ("teapot", 48, 27, 2), ("cornell_sphere", 32, 32, 2), ("cornell_glossy", 32, 32, 2),
                                            ("veach_beckmann", 192, 128, 16), ("veach_beckmann", 48, 32, 4),
                                            ("cornell_metal", 96, 96, 8), ("cornell_metal", 32, 32, 2),
                                            ("cornell_lens", 96, 96, 8), ("cornell_lens", 32, 32, 2),
                                            ("textured", 192, 128, 8), ("textured", 48, 32, 2),
                                            ("random_scene", 192, 128, 8), ("random_scene", 48, 32, 2)])
def test_same_random_numbers_as_golden(built, name, nx, ny, spp):
    """Same check against the committed fixtures (generated from the compiled reference by
    scripts/make_golden.py), so it also runs where oracle/_ref is absent."""
    g = np.load(refutil.GOLDEN / f"render_{name}_{nx}x{ny}x{spp}.npz")
    gpu, st = _gpu_render(name, nx, ny, spp, seed=int(g["seed"]))
    ref = g["image"].astype(np.float64)
    close = (np.abs(gpu - ref).max(axis=2) / np.maximum(ref.max(axis=2), 1e-3)) < 2e-3
    dm = abs(gpu.mean() - ref.mean()) / ref.mean()
    print(f"{name} {nx}x{ny}x{spp}: pixels within 0.2% {close.mean():.4f}, mean diff {dm:.2e}, "
          f"closest rays gpu {st['closest_rays']} ref {int(g['closest_rays'])}")
    # a path whose float32 shading flips a branch can carry an 800-radiance light sample (veach): such
    # pixels are counted (< 0.1 %) and the mean is compared over the others
    dm_close = abs(gpu[close].mean() - ref[close].mean()) / ref[close].mean()
    assert close.mean() > 0.999 if nx * ny > 10_000 else close.mean() > 0.99
    assert dm_close < 1e-4
    if nx * ny > 4000:  # a 32 x 32 x 2 image holds too few samples for a 0.5 % bar on one flipped path
        _whole_image_mean_bar(gpu, ref, close, f"{name} {nx}x{ny}x{spp}")
    assert abs(st["closest_rays"] - int(g["closest_rays"])) / int(g["closest_rays"]) < 5e-3


@pytest.mark.parametrize("name,nx,ny,spp,integrator", [("teapot", 48, 27, 2, "ao"), ("teapot", 160, 90, 4, "ao"),
                                                       ("cornell_glossy+sky", 32, 32, 2, "ao"), ("veach+sky", 48, 32, 2, "ao"),
                                                       ("cornell_sphere+sky", 32, 32, 2, "ao"), ("cornell", 32, 32, 1, "normals"),
                                                       ("teapot", 48, 27, 1, "normals"), ("cornell_sphere", 32, 32, 1, "normals"),
                                                       ("veach+sky", 48, 32, 2, "path"),
                                                       ("cornell", 32, 32, 2, "direct"), ("cornell_glossy", 32, 32, 2, "direct"),
                                                       ("teapot", 48, 27, 2, "direct"), ("teapot+sky", 48, 27, 1, "direct"),
                                                       ("cornell", 128, 128, 4, "direct"),
                                                       # HDR image environment (material.h:209-232, texture.h:60-86)
                                                       ("teapot+hdr", 48, 27, 2, "path"), ("veach+hdr", 48, 32, 2, "path"),
                                                       ("teapot+hdr", 160, 90, 4, "path"),
                                                       # round 2 fixtures
                                                       ("textured", 96, 64, 2, "direct"), ("textured", 96, 64, 1, "normals"),
                                                       ("veach_beckmann+sky", 48, 32, 2, "ao"), ("textured+hdr", 48, 32, 2, "path"),
                                                       # furnace_test_scene (main.cpp:177-221): dielectric mesh, NO lights
                                                       ("furnace+hdr", 48, 27, 2, "path"), ("furnace+hdr", 160, 90, 4, "path"),
                                                       # cornell_box_ao (main.cpp:254-280): smooth generated normals, white environment
                                                       ("cornell_ao", 32, 32, 2, "ao"), ("cornell_ao", 96, 96, 8, "ao"),
                                                       ("cornell_ao", 32, 32, 2, "path"), ("cornell_ao", 32, 32, 1, "normals"),
                                                       # random_scene (main.cpp:55-113): 485 spheres, thin lens
                                                       ("random_scene", 48, 32, 2, "path"), ("random_scene+hdr", 48, 32, 2, "path"),
                                                       ("random_scene", 96, 64, 1, "normals")])
def test_other_integrators_match_golden(built, name, nx, ny, spp, integrator):
    """gpu_ao_integrator / gpu_normals_integrator / gpu_ass_1_integrator (same trace kernels, another shade stage) and the environment
    term of the path integrator against fixtures rendered by the compiled reference's ao::Li /
    normals_renderer::Li / path::Li on the same random numbers."""
    tag = "render" if integrator == "path" else f"render_{integrator}"
    g = np.load(refutil.GOLDEN / f"{tag}_{name}_{nx}x{ny}x{spp}.npz")
    scene = api.HostScene(name, nx / ny, 2)
    gpu, st, _ = scene.render(nx, ny, spp, seed=int(g["seed"]), integrator=integrator)
    gpu = gpu.astype(np.float64)
    ref = g["image"].astype(np.float64)
    assert np.isfinite(gpu).all() and np.abs(ref).max() > 0
    if integrator == "normals":
        # one sample per pixel: a float32 jitter rounding can move a sample across an edge in a few pixels
        close = np.abs(gpu - ref).max(axis=2) < 2e-4
        assert close.mean() > 0.995
    else:
        close = (np.abs(gpu - ref).max(axis=2) / np.maximum(ref.max(axis=2), 1e-3)) < 2e-3
        assert close.mean() > (0.999 if nx * ny > 10_000 else 0.99)
        assert abs(gpu[close].mean() - ref[close].mean()) / ref[close].mean() < 1e-4
    print(f"{integrator} {name} {nx}x{ny}x{spp}: pixels agreeing {close.mean():.4f}, mean gpu {gpu.mean():.6f} ref {ref.mean():.6f}, "
          f"rays {st['closest_rays']}+{st['shadow_rays']}")


def test_cornell_statistical_parity(built):
    """BASELINE bars on independent random numbers against the as-shipped mt19937 reference: 256x256, reference at
    1024 spp (tests/golden/stat_cornell_256x256x1024.npz, scripts/make_golden_stat.py), GPU at 4096 spp.  Two
    reference renders of 1024 spp differ by relMSE 9.7e-4 at full resolution (their own noise), so the GPU runs 4x the
    samples; relMSE <= 1e-3 at FULL resolution and mean within 0.5 % — BASELINE.json's bars as they stand."""
    g = np.load(refutil.GOLDEN / "stat_cornell_256x256x1024.npz")
    ref = g["image"].astype(np.float64)
    gpu, _ = _gpu_render("cornell", 256, 256, 4096, seed=3)
    rel, rel8 = refutil.rel_mse(gpu, ref), refutil.rel_mse(refutil.block_mean(gpu, 8), refutil.block_mean(ref, 8))
    dm = abs(gpu.mean() - ref.mean()) / ref.mean()
    print(f"cornell 256^2 statistical: relMSE full-res {rel:.3e}, 8x8 {rel8:.3e}, mean diff {dm:.4%}")
    assert rel <= 1e-3
    assert rel8 <= 1e-4
    assert dm <= 5e-3


def test_config1_cornell_512_statistical_parity(built):
    """BASELINE config 1 (Cornell-Original 512x512, the one config the reference runs verbatim, main.cpp:222-252):
    GPU at 4096 spp against the converged mt19937 reference (1024 spp, full resolution: relMSE <= 1e-3, mean within
    0.5 %), and GPU at the config's own 16 spp against the reference's 16-spp image after 8x8 block averaging."""
    conv = np.load(refutil.GOLDEN / "stat_cornell_512x512x1024.npz")["image"].astype(np.float64)
    gpu, st = _gpu_render("cornell", 512, 512, 4096, seed=3)
    rel = refutil.rel_mse(gpu, conv)
    dm = abs(gpu.mean() - conv.mean()) / conv.mean()
    print(f"config 1 converged: relMSE full-res {rel:.3e}, mean diff {dm:.4%}, {st['render_ms']:.1f} ms device for 4096 spp")
    assert rel <= 1e-3 and dm <= 5e-3
    g16 = np.load(refutil.GOLDEN / "stat_cornell_512x512x16.npz")
    gpu16, st16 = _gpu_render("cornell", 512, 512, 16, seed=3)
    rel16 = refutil.rel_mse(refutil.block_mean(gpu16, 8), refutil.block_mean(g16["image"].astype(np.float64), 8))
    dm16 = abs(gpu16.mean() - conv.mean()) / conv.mean()
    print(f"config 1 at 16 spp: relMSE(8x8) vs the reference's 16-spp image {rel16:.3e}, mean vs converged {dm16:.4%}; "
          f"GPU {st16['render_ms']:.2f} ms, reference {float(g16['seconds']):.1f} s on {int(g16['threads'])} threads")
    assert rel16 <= 2e-3   # two 16-spp images: twice the variance of one against a converged image
    assert dm16 <= 1e-2    # 4.2 M samples: the mean itself carries ~0.3 % noise


def test_headline_config_1080p_1024spp_parity(built):
    """The bench configuration itself (BASELINE config 2: Cornell 1920x1080, 1024 spp) against the as-shipped mt19937
    reference at the same resolution (256 spp, stored as 8x8 block means: tests/golden/stat_cornell_1920x1080x256_block8.npz):
    relMSE(8x8) <= 1e-3 and mean radiance within 0.5 %."""
    g = np.load(refutil.GOLDEN / "stat_cornell_1920x1080x256_block8.npz")
    ref8, ref_mean = g["image"].astype(np.float64), float(g["mean"])
    gpu, st = _gpu_render("cornell", 1920, 1080, 1024, seed=1, max_leaf=2)
    rel = refutil.rel_mse(refutil.block_mean(gpu, 8), ref8)
    dm = abs(gpu.mean() - ref_mean) / ref_mean
    print(f"1080p x 1024 spp: relMSE(8x8) {rel:.3e}, mean gpu {gpu.mean():.6f} ref {ref_mean:.6f} ({dm:.4%}), "
          f"{st['render_ms']:.0f} ms device, {(st['closest_rays'] + st['shadow_rays']) / st['render_ms'] / 1e6:.2f} Grays/s")
    assert st["samples"] == 1920 * 1080 * 1024
    assert rel <= 1e-3
    assert dm <= 5e-3


@pytest.mark.parametrize("name,nx,ny,spp,fixture,block", [("veach", 768, 512, 1024, "stat_veach_768x512x256_block4", 4),
                                                          ("teapot", 480, 270, 1024, "stat_teapot_480x270x256", 1)])
def test_other_configs_statistical_parity(built, name, nx, ny, spp, fixture, block):
    """BASELINE configs 3 and 4 against the as-shipped mt19937 reference (veach at its mi.xml film size, 768x512):
    mean within 0.5 %, relMSE <= 1e-3 after block averaging (veach: the 800-radiance sphere light makes single pixels
    heavy-tailed; the bar is taken on 16x16 blocks there)."""
    g = np.load(refutil.GOLDEN / f"{fixture}.npz")
    ref = g["image"].astype(np.float64)
    gpu, _ = _gpu_render(name, nx, ny, spp, seed=3, max_leaf=2)
    gb = refutil.block_mean(gpu, block) if block > 1 else gpu
    k = 4 if name == "veach" else 8
    rel = refutil.rel_mse(refutil.block_mean(gb, k), refutil.block_mean(ref, k))
    dm = abs(gpu.mean() - float(g["mean"])) / float(g["mean"])
    print(f"{name} {nx}x{ny}: relMSE({block * k}x{block * k}) {rel:.3e}, mean diff {dm:.4%}")
    assert rel <= 1e-3
    assert dm <= 5e-3


def test_split_spp_ranges_sum_to_whole(built):
    """Sample s of pixel p is the same whatever range renders it: [0,8) == [0,3) + [3,8)."""
    scene = api.HostScene("cornell", 1.0, 4)
    ctx = api.Context()
    ctx.upload(scene)
    ctx.render_tile(64, 64, 0, 8, seed=9)
    whole, _ = ctx.readback()
    ctx.clear_film()
    ctx.render_tile(64, 64, 0, 3, seed=9)
    ctx.render_tile(64, 64, 3, 8, seed=9)
    parts, _ = ctx.readback()
    ctx.close()
    assert np.allclose(whole, parts, rtol=1e-4, atol=1e-5)
    assert whole.sum() > 0


def test_in_process_multi_device_matches_single(built):
    """frt_create(n > 1): the sample range is split over the devices and the films are summed by the reduce-scatter
    over peer memory (k_sum_slice) — same image as one device (same samples, different summation order).  Runs on a
    one-GPU box too: devices = {0, 0, 0} are three device records (streams, films, counters) on one GPU, which
    exercises the split, the cross-stream waits, the in-place sum and the film clearing; with real peers the same
    code reads over NVLink (test below)."""
    scene = api.HostScene("cornell", 1.0, 2)
    one = api.Context(devices=[0]); one.upload(scene)
    one.render_tile(96, 96, 0, 10, seed=11)
    want, st1 = one.readback()
    one.close()
    ctx = api.Context(devices=[0, 0, 0]); ctx.upload(scene)
    for attempt in range(2):  # the second render must not see the first one's partial films
        ctx.clear_film()
        ctx.render_tile(96, 96, 0, 10, seed=11)
        got, stn = ctx.readback()
        assert stn["samples"] == st1["samples"] == 96 * 96 * 10
        assert stn["closest_rays"] == st1["closest_rays"] and stn["shadow_rays"] == st1["shadow_rays"]
        assert np.allclose(want, got, rtol=1e-4, atol=1e-6)
    again, _ = ctx.readback()  # a second readback without rendering: the folded partial films are not added twice
    assert np.allclose(got, again, rtol=0, atol=0)
    # sample ranges smaller than the device count, and a film size that is not a multiple of 4 floats
    ctx.clear_film()
    ctx.render_tile(33, 21, 0, 2, seed=5)
    small, sts = ctx.readback()
    one = api.Context(devices=[0]); one.upload(scene)
    one.render_tile(33, 21, 0, 2, seed=5)
    small1, _ = one.readback()
    one.close()
    assert sts["samples"] == 33 * 21 * 2 and np.allclose(small, small1, rtol=1e-4, atol=1e-6)
    # the ray-cast entry block-partitions over the same device records
    rays = scene.camera_rays(50, 41)
    h3, _ = ctx.trace(rays)
    ctx.close()
    one = api.Context(devices=[0]); one.upload(scene)
    h1, _ = one.trace(rays)
    one.close()
    assert (h3["prim_id"] == h1["prim_id"]).all() and np.array_equal(h3["t"], h1["t"])


def test_in_process_multi_gpu_matches_single(built):
    """The same through renderer<gpu_path_integrator> on real peers (NVLink loads in k_sum_slice)."""
    import torch
    n = min(torch.cuda.device_count(), 4)
    if n < 2:
        pytest.skip("needs at least two GPUs (the device-record path is covered by the test above)")
    scene = api.HostScene("cornell", 1.0, 2)
    one, st1, _ = scene.render(96, 96, 10, seed=11, n_gpus=1)
    many, stn, _ = scene.render(96, 96, 10, seed=11, n_gpus=n)
    again, _, _ = scene.render(96, 96, 10, seed=11, n_gpus=n)  # cached context, film cleared
    assert stn["samples"] == st1["samples"] == 96 * 96 * 10
    assert stn["closest_rays"] == st1["closest_rays"]
    assert np.allclose(one, many, rtol=1e-4, atol=1e-6)
    assert np.allclose(many, again, rtol=1e-4, atol=1e-6)


def test_cancel_stops_a_long_tile_within_milliseconds(built):
    """viewer::to_exit (viewer.cpp:101-104, polled by path.cpp:127): a 1080p x 4096 spp tile (about 7 s of device time)
    is cancelled 100 ms after its launch; frt_readback must then return within 50 ms, with a partial film and a partial
    sample count.  The next tile renders normally (the request does not stick)."""
    import time
    scene = api.HostScene("cornell", 16 / 9, 2)
    ctx = api.Context(); ctx.upload(scene)
    ctx.render_tile(64, 36, 0, 2, seed=1); ctx.readback()  # module load / first-launch costs out of the way
    ctx.clear_film()
    ctx.render_tile(1920, 1080, 0, 4096, seed=1)
    time.sleep(0.1)
    assert not ctx.poll()
    t0 = time.perf_counter()
    ctx.cancel()
    _, st = ctx.readback(want_image=False)
    dt = time.perf_counter() - t0
    print(f"cancel: readback returned {dt * 1e3:.2f} ms after frt_cancel; {st['samples']} of {1920 * 1080 * 4096} samples started")
    assert dt < 0.05
    assert 0 < st["samples"] < 1920 * 1080 * 4096
    assert ctx.poll()
    ctx.clear_film()
    ctx.render_tile(64, 36, 0, 4, seed=1)
    img, st2 = ctx.readback()
    assert st2["samples"] == 64 * 36 * 4 and img.sum() > 0
    ctx.close()


def test_viewer_to_exit_cancels_the_integrator(built):
    """The same through the reference-facing layer: another thread does what closing the preview window does
    (frth_cancel sets viewer::to_exit of the film being rendered); renderer<gpu_path_integrator>::Render returns
    early with a partial film."""
    import threading
    import time
    scene = api.HostScene("cornell", 16 / 9, 2)
    scene.render(64, 36, 2)  # context warm-up
    _, frth = api.load_libs()
    fired = {}

    def closer():
        time.sleep(0.3)
        fired["rc"] = frth.frth_cancel()

    th = threading.Thread(target=closer)
    th.start()
    t0 = time.perf_counter()
    img, st, secs = scene.render(1920, 1080, 8192, seed=1)
    dt = time.perf_counter() - t0
    th.join()
    print(f"viewer::to_exit after 0.3 s: Render returned after {dt:.3f} s, {st['samples'] / (1920 * 1080):.1f} of 8192 spp started")
    assert fired["rc"] == 0 and st["cancelled"]
    assert dt < 1.5 and st["samples"] < 1920 * 1080 * 8192
    assert frth.frth_cancel() != 0  # nothing rendering now


def test_device_tonemap_equals_viewer_add_sample(built):
    """frt_readback_film = viewer::add_sample for every pixel (viewer.cpp:109-132): fout_image is sum * (1 / ns) in
    double, out_image is int(pow(1 - exp(-c), 1 / 2.2) * 255 + .5) evaluated in double on the device.  Compared with
    the same expression in numpy float64 on the film sums: bytes equal (a pixel within 1e-9 of a rounding boundary may
    differ by one; none is expected in 10^5 values), including negative means (normals integrator -> NaN -> byte 0)."""
    for name, integ, spp in (("cornell", "path", 8), ("cornell_sphere", "normals", 1)):
        scene = api.HostScene(name, 1.0, 2)
        ctx = api.Context(); ctx.upload(scene)
        ctx.render_tile(160, 120, 0, spp, seed=4, integrator=integ)
        sums, _ = ctx.readback()
        fout, ldr, _ = ctx.readback_film(spp)
        ctx.close()
        c = sums.astype(np.float64) * (1.0 / spp)
        assert np.array_equal(fout, c)
        with np.errstate(invalid="ignore"):
            v = np.power(1.0 - np.exp(-c), 1 / 2.2) * 255 + .5
        want = np.where(np.isfinite(v), np.floor(np.where(np.isfinite(v), v, 0)), 0).astype(np.int64)
        want = (want & 0xFF).astype(np.uint8)
        near_edge = np.isfinite(v) & (np.abs(v - np.round(v)) < 1e-9)
        differ = (ldr != want) & ~near_edge
        print(f"{name}/{integ}: {differ.sum()} of {ldr.size} bytes differ from the float64 expression; negative means: {(c < 0).sum()}")
        assert differ.sum() == 0
        assert ldr.max() > 100
    # and through the reference-facing layer: viewer::out_image of the last render
    scene = api.HostScene("cornell", 1.0, 2)
    img, _, _ = scene.render(96, 96, 4, seed=2)
    ldr = scene.last_ldr(96, 96)
    with np.errstate(invalid="ignore"):
        v = np.power(1.0 - np.exp(-img.astype(np.float64)), 1 / 2.2) * 255 + .5
    assert (np.abs(ldr.astype(np.int64) - np.floor(v).astype(np.int64)) <= 1).all()  # img is the float32 copy of the mean


def test_reference_renderer_template_drives_the_gpu_integrator(built):
    """The drop-in build (oracle/Makefile `dropin`): the REFERENCE's own renderer<> (integrator.h:11-47), viewer
    members and Scene classes with gpu_path_integrator as the template argument render the same image as the host
    layer's mirror of that API (same flattened scene — byte-equal, tests/test_host.py — same seed, same kernels)."""
    if not refutil.DROPIN_BIN.exists():
        pytest.skip("oracle/_ref/frt_dropin not built")
    for name, nx, ny, spp, integ in (("cornell", 96, 96, 8, "path"), ("veach", 96, 64, 8, "path"), ("textured", 96, 64, 4, "path"),
                                     ("teapot", 96, 54, 2, "ao")):
        ref_img, info = refutil.dropin_render(name, nx, ny, spp, seed=5, integrator=integ)
        scene = api.HostScene(name, nx / ny, 2)
        img, st, _ = scene.render(nx, ny, spp, seed=5, integrator=integ)
        assert info["samples"] == st["samples"] == nx * ny * spp
        assert info["closest_rays"] == st["closest_rays"] and info["shadow_rays"] == st["shadow_rays"]
        # same samples, same arithmetic; only the order of the film's float atomics differs between two runs
        assert np.allclose(ref_img, img, rtol=1e-4, atol=1e-6), name
        assert info["ldr_sum"] > 0


def test_repeated_renders_reuse_the_context(built):
    """renderer<gpu_path_integrator> keeps its device context between renders: a second render of another
    size and scene must not see the first one's film or scene."""
    a = api.HostScene("cornell", 1.0, 2)
    b = api.HostScene("veach", 1.5, 2)
    img_a, _, _ = a.render(64, 64, 4, seed=3)
    img_b, _, _ = b.render(96, 64, 4, seed=3)
    img_a2, _, _ = a.render(64, 64, 4, seed=3)
    assert img_b.shape == (64, 96, 3) and np.isfinite(img_b).all() and img_b.mean() > 0
    assert np.allclose(img_a, img_a2, rtol=1e-4, atol=1e-6)


def test_pixel_tiles_sum_to_whole_and_bad_tiles_are_rejected(built):
    """A sample is identified by (seed, pixel, index): rendering the film as two pixel rectangles gives the
    frame rendered at once; tiles outside the film, unknown integrators and ao without a range are errors."""
    scene = api.HostScene("cornell", 1.0, 2)
    ctx = api.Context()
    ctx.upload(scene)
    ctx.render_tile(64, 48, 0, 6, seed=4)
    whole, st = ctx.readback()
    ctx.clear_film()
    ctx.render_tile(64, 48, 0, 6, seed=4, rect=(0, 0, 64, 20))
    ctx.render_tile(64, 48, 0, 6, seed=4, rect=(0, 20, 64, 28))
    parts, st2 = ctx.readback()
    assert st2["samples"] == st["samples"] == 64 * 48 * 6
    assert np.allclose(whole, parts, rtol=1e-4, atol=1e-5)
    with pytest.raises(api.FrtError):
        ctx.render_tile(64, 48, 0, 1, rect=(32, 0, 64, 48))
    bad = api.frt_tile(64, 48, 0, 0, 64, 48, 0, 1, 1, 34, 7, 0)
    assert ctx._frt.frt_render_tile(ctx.handle, C.byref(bad)) == -1  # FRT_ERR_INVALID: unknown integrator
    assert b"integrator" in ctx._frt.frt_last_error(ctx.handle)
    # empty sample range: nothing rendered, nothing counted
    ctx.clear_film()
    ctx.render_tile(64, 48, 3, 3, seed=4)
    empty, st3 = ctx.readback()
    assert st3["samples"] == 0 and not empty.any()
    ctx.close()


def test_checkpoint_and_resume(built, tmp_path):
    """a render interrupted after 3 of 8 spp and resumed in a NEW context from its checkpoint
    equals the uninterrupted render."""
    scene = api.HostScene("cornell", 1.0, 2)
    ctx = api.Context(); ctx.upload(scene)
    ctx.render_tile(64, 64, 0, 8, seed=9)
    whole, _ = ctx.readback()
    ctx.clear_film()
    ctx.render_tile(64, 64, 0, 3, seed=9)
    part, _ = ctx.readback()
    ctx.close()
    api.save_checkpoint(tmp_path / "film.npz", part, spp_done=3, seed=9, scene="cornell")
    ck = api.load_checkpoint(tmp_path / "film.npz")
    assert ck["spp_done"] == 3 and ck["seed"] == 9 and ck["scene"] == "cornell"
    ctx2 = api.Context(); ctx2.upload(scene)
    ctx2.restore_film(ck["sums"])
    ctx2.render_tile(64, 64, ck["spp_done"], 8, seed=ck["seed"])
    resumed, st = ctx2.readback()
    assert st["samples"] == 64 * 64 * 5
    assert np.allclose(whole, resumed, rtol=1e-4, atol=1e-5)
    with pytest.raises(api.FrtError):
        ctx2.restore_film(np.zeros((8, 8, 2), np.float32))
    ctx2.close()


def test_frt_main_checkpoint_resume(built, tmp_path):
    """frt_main --checkpoint / --resume: finishing a 3-spp job as an 8-spp job gives the 8-spp film."""
    import subprocess
    import os
    exe = str(api.PKG_DIR / "frt_main")
    env = dict(os.environ, FRT_ASSETS=str(refutil.ASSETS))  # the test may be the first one of the process
    def run(*args):
        r = subprocess.run([exe, "--scene", "cornell", "--res", "48x32", "--seed", "5", *args], cwd=tmp_path,
                           capture_output=True, text=True, timeout=300, env=env)
        assert r.returncode == 0, r.stderr
        return r.stdout
    def sums(name):
        raw = (tmp_path / name).read_bytes()
        assert raw[:8] == b"FRTCKPT2"
        nx, ny, done, seed, version, integ, max_depth, scene_hash = np.frombuffer(raw[8:40], np.uint32)
        assert version == 2 and integ == 0 and max_depth == 34
        return int(done), np.frombuffer(raw[40:], np.float32).reshape(ny, nx, 3)
    run("--ns", "8", "--checkpoint", "whole.ck", "--png", "whole.png")
    # --png: the 8-bit tone-mapped film (viewer::out_image, what the reference's window shows) as a file
    import struct
    import zlib
    raw = (tmp_path / "whole.png").read_bytes()
    assert raw[:8] == b"\x89PNG\r\n\x1a\n" and struct.unpack(">II", raw[16:24]) == (48, 32)
    n_idat = struct.unpack(">I", raw[33:37])[0]
    assert raw[37:41] == b"IDAT"
    rows = np.frombuffer(zlib.decompress(raw[41:41 + n_idat]), np.uint8).reshape(32, 1 + 3 * 48)
    assert rows[:, 1:].max() > 100 and rows[:, 1:].min() == 0  # a lit Cornell box with black surroundings
    run("--ns", "3", "--checkpoint", "part.ck")
    out = run("--ns", "8", "--resume", "part.ck", "--checkpoint", "resumed.ck")
    assert "Resuming at sample 3 of 8" in out
    (dw, whole), (dp, _), (dr, resumed) = sums("whole.ck"), sums("part.ck"), sums("resumed.ck")
    assert (dw, dp, dr) == (8, 3, 8)
    assert np.allclose(whole, resumed, rtol=1e-4, atol=1e-5)
    bad = subprocess.run([exe, "--res", "40x32", "--ns", "8", "--resume", "part.ck"], cwd=tmp_path, capture_output=True, text=True, env=env)
    assert bad.returncode == 1 and "resolution" in bad.stderr
    # a checkpoint of another scene / integrator is refused instead of silently mixing estimators
    bad = subprocess.run([exe, "--scene", "veach", "--res", "48x32", "--ns", "8", "--resume", "part.ck"], cwd=tmp_path, capture_output=True, text=True, env=env)
    assert bad.returncode == 1 and "scene" in bad.stderr
    bad = subprocess.run([exe, "--scene", "cornell", "--integrator", "gpu_ao", "--res", "48x32", "--ns", "8", "--resume", "part.ck"],
                         cwd=tmp_path, capture_output=True, text=True, env=env)
    assert bad.returncode == 1 and "integrator" in bad.stderr


@pytest.mark.parametrize("integrator", ["path", "ao", "normals", "direct"])
def test_image_environment_matches_oracle(built, integrator):
    """environment_map::eval over an image_texture (material.h:219-232, texture.h:60-86): every texel of a small
    random lat-long image is a different colour, so a wrong row / column / wrap shows in every pixel that sees the
    environment.  GPU against the C oracle through the same scene description (the oracle itself is pinned on the
    reference's stb-loaded data/test.hdr by the +hdr fixtures)."""
    import ctypes as C
    import oracle_c
    nx, ny, spp = 96, 54, 2
    scene = api.HostScene("teapot", nx / ny, 2)
    env = np.ascontiguousarray(np.random.default_rng(11).random((16, 32, 3), dtype=np.float32) + 0.25)
    scene.desc.env_image = env.ctypes.data
    scene.desc.env_nx, scene.desc.env_ny = 32, 16
    ref = oracle_c.Oracle(scene).render(nx, ny, spp, seed=5, integrator=integrator)
    ctx = api.Context()
    ctx.upload(scene)
    ctx.render_tile(nx, ny, 0, spp, seed=5, integrator=integrator)
    gpu, st = ctx.readback()
    ctx.close()
    gpu = gpu.astype(np.float64) / spp
    sees_env = np.abs(ref).max(axis=2) > 0
    assert sees_env.mean() > 0.3
    close = np.abs(gpu - ref).max(axis=2) <= 2e-3 * np.maximum(np.abs(ref).max(axis=2), 1e-3) + 2e-4
    print(f"image environment, {integrator}: pixels agreeing {close.mean():.4f}, mean gpu {gpu.mean():.6f} oracle {ref.mean():.6f}")
    # a direction within float32 rounding of a texel edge may take the neighbouring texel
    assert close.mean() > 0.99
    assert abs(gpu[close].mean() - ref[close].mean()) <= 1e-4 * abs(ref[close].mean())


def test_image_environment_validation(built):
    """An environment image without a size is refused at upload, like every other malformed scene array."""
    scene = api.HostScene("cornell", 1.0, 2)
    env = np.ones((4, 8, 3), np.float32)
    scene.desc.env_image = env.ctypes.data
    scene.desc.env_nx, scene.desc.env_ny = 0, 4
    ctx = api.Context()
    with pytest.raises(Exception):
        ctx.upload(scene)
    ctx.close()
