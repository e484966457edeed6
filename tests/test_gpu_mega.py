"""RT_KERNEL_MEGA (persistent megakernel over the rebuilt BVH) against the oracle.

Hit tests: the rebuilt structure changes only the ORDER in which candidates are met, so hit distances must be the
reference's bits (bar: <= 1 ulp; observed: 0) with identical object ids.  Images: same seeds, same spp, compared by
per-channel RMSE against the oracle, with the tolerance calibrated by the oracle's own noise floor."""
from __future__ import annotations

import numpy as np
import pytest

import raygen
from conftest import same_bits, ulp_distance

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def gpu_scene(rtc, scene_factory):
    cache = {}

    def make(scene_id, width=48, spp=4, **kw):
        key = (scene_id, width, spp, tuple(sorted(kw.items())))
        if key not in cache:
            sess, desc = scene_factory(scene_id, width, spp, **kw)
            cache[key] = (sess, desc, rtc.Scene(desc))
        return cache[key]

    yield make
    for _, _, scene in cache.values():
        scene.close()


def _batch(sess, desc, golden_file, n=200_000):
    o2, d2 = raygen.camera_rays(sess.camera, n, seed=201)
    o3, d3 = raygen.secondary_rays(desc, n, seed=202)
    origins = np.concatenate([golden_file["rays_o"], o2, o3])
    dirs = np.concatenate([golden_file["rays_d"], d2, d3])
    # d = (0,0,0) is not a ray: the reference's sphere test then returns t = 0/0 = NaN as a "hit" of whatever it
    # visits first; only the reference-order kernel reproduces that (test_gpu_exact keeps those rays)
    real = (dirs != 0).any(axis=1)
    return origins[real], dirs[real]


@pytest.mark.parametrize("scene_id", [0, 1, 2, 3, 4, 5, 6])
def test_hit_distance_and_id_parity(rtc, oracle, golden, gpu_scene, scene_id):
    sess, desc, scene = gpu_scene(scene_id)
    origins, dirs = _batch(sess, desc, golden(scene_id))
    want = oracle.hit_batch(desc, origins, dirs, math=oracle.MATH_PORTABLE)
    got = scene.hit_batch(origins, dirs, kernel="mega")
    assert np.array_equal(got["node"] != rtc.abi.RT_NONE, want["node"] != rtc.abi.RT_NONE)
    assert ulp_distance(got["t"], want["t"]).max() <= 1
    assert same_bits(got["t"], want["t"]).all()
    assert np.array_equal(got["node"], want["node"])
    hit = want["node"] != rtc.abi.RT_NONE
    assert same_bits(got["p"][hit], want["p"][hit]).all()
    assert same_bits(got["normal"][hit], want["normal"][hit]).all()
    assert np.array_equal(got["front"][hit], want["front"][hit])
    assert np.array_equal(got["material"], want["material"])
    assert np.nanmax(np.abs(got["uv"][hit] - want["uv"][hit]), initial=0.0) < 1e-5


def test_hit_parity_second_bounce_on_bvh_scenes(rtc, oracle, gpu_scene):
    for scene_id in (1, 6):
        sess, desc, scene = gpu_scene(scene_id)
        o, d = raygen.camera_rays(sess.camera, 100_000, seed=9)
        first = scene.hit_batch(o, d, kernel="mega")
        ok = (first["node"] != rtc.abi.RT_NONE) & np.isfinite(first["p"]).all(axis=1)
        o2, d2 = raygen.surface_rays(desc, first["p"], ok, 200_000, seed=10)
        want = oracle.hit_batch(desc, o2, d2, math=oracle.MATH_PORTABLE)
        got = scene.hit_batch(o2, d2, kernel="mega")
        assert same_bits(got["t"], want["t"]).all()
        assert np.array_equal(got["node"], want["node"])


def test_hit_parity_with_media(rtc, oracle, golden, gpu_scene):
    """Scene 7 (two ConstantMedium items in the world list).  The rebuilt structure traces all surfaces first and
    then reproduces the reference's in-list evaluation of the media from that (trace_media, rtc_trace.cuh): the same
    hit, the same hit distance bits and the SAME NUMBER OF RANDOM DRAWS on every ray -- so ids, t and the advanced
    PCG32 states are compared for equality, media included (/root/reference/src/hittable.c:74-88,392-423)."""
    sess, desc, scene = gpu_scene(7)
    origins, dirs = _batch(sess, desc, golden(7), n=150_000)
    rng = raygen.rng_states(len(origins), seed=203)
    want = oracle.hit_batch(desc, origins, dirs, rng=rng, math=oracle.MATH_PORTABLE)
    got = scene.hit_batch(origins, dirs, rng=rng, kernel="mega")
    is_medium = desc.nodes["kind"] == rtc.abi.NODE_MEDIUM
    med_w = (want["node"] != rtc.abi.RT_NONE) & is_medium[np.minimum(want["node"], len(is_medium) - 1)]
    assert 0.02 < med_w.mean() < 0.9  # the batch does exercise the media
    drew = want["rng"][:, 0] != rng[:, 0]
    assert drew.mean() > 0.5          # ... and most rays draw (the r = 5000 fog ball encloses the scene)
    assert np.array_equal(got["rng"], want["rng"])
    assert np.array_equal(got["node"], want["node"])
    assert same_bits(got["t"], want["t"]).all()
    hit = want["node"] != rtc.abi.RT_NONE
    assert same_bits(got["p"][hit], want["p"][hit]).all()
    assert np.array_equal(got["material"], want["material"])
    # rays whose closest surface is a LATER list item than a medium they also cross: the case the restricted
    # queries of trace_media exist for -- present in the batch, not just in theory
    node = want["node"].astype(np.int64)
    med_ids = np.flatnonzero(is_medium)
    later_surface = hit & ~med_w & (node > med_ids.min())
    assert (later_surface & drew).sum() > 100


def _rmse(a, b):
    a, b = np.clip(a, 0, 1), np.clip(b, 0, 1)
    return np.sqrt(((a - b) ** 2).mean(axis=(0, 1)))


@pytest.mark.parametrize("scene_id", range(8))
def test_render_rmse_against_oracle(rtc, oracle, gpu_scene, scene_id):
    """Tolerance (stated, per channel, on radiance clipped to [0,1]):
         RMSE(mega, oracle) <= 0.5 * RMSE(oracle stream A, oracle stream B)   -- far inside the Monte-Carlo noise
         |mean(mega) - mean(oracle)| <= 2/255
    Both sides use the per-sample seeds, so most paths are literally the same path; they only fork where fast
    intrinsics flip a decision."""
    width, spp = (64, 16) if scene_id != 7 else (48, 16)
    sess, desc, scene = gpu_scene(scene_id, width, spp)
    cam = sess.camera
    _, rad_o, st_o = oracle.render(desc, cam, rtc.make_opts(kernel="exact", rng="sample"), oracle.MATH_PORTABLE)
    _, rad_p, _ = oracle.render(desc, cam, rtc.make_opts(kernel="exact", rng="pixel"), oracle.MATH_PORTABLE)
    _, rad_m, st_m = scene.render(cam, rtc.make_opts(kernel="mega", rng="sample"))
    img_o, img_p, img_m = rad_o / spp, rad_p / spp, rad_m / spp
    ok = np.isfinite(img_o).all(axis=2) & np.isfinite(img_m).all(axis=2) & np.isfinite(img_p).all(axis=2)
    assert ok.mean() > 0.99
    img_o, img_p, img_m = (np.where(ok[..., None], x, 0.0) for x in (img_o, img_p, img_m))
    floor = _rmse(img_o, img_p)
    got = _rmse(img_m, img_o)
    assert (got <= 0.5 * floor + 1e-4).all(), (got, floor)
    bias = np.abs(np.clip(img_m, 0, 1).mean(axis=(0, 1)) - np.clip(img_o, 0, 1).mean(axis=(0, 1)))
    assert (bias <= 2.0 / 255.0).all(), bias
    assert st_m["samples"] == st_o["samples"]
    assert abs(st_m["segments"] - st_o["segments"]) <= 0.02 * st_o["segments"]


@pytest.mark.parametrize("scene_id", range(8))
def test_wavefront_render_rmse_against_oracle(rtc, oracle, gpu_scene, scene_id):
    """RT_KERNEL_WAVE (generate / extend+classify / shade-per-material queues) under the same bar as the megakernel,
    plus: it traces the same number of camera samples and (within fast-math forks) the same segments as the
    megakernel, and the two images agree far inside the noise."""
    width, spp = (64, 16) if scene_id != 7 else (48, 16)
    sess, desc, scene = gpu_scene(scene_id, width, spp)
    cam = sess.camera
    _, rad_o, st_o = oracle.render(desc, cam, rtc.make_opts(kernel="exact", rng="sample"), oracle.MATH_PORTABLE)
    _, rad_p, _ = oracle.render(desc, cam, rtc.make_opts(kernel="exact", rng="pixel"), oracle.MATH_PORTABLE)
    _, rad_w, st_w = scene.render(cam, rtc.make_opts(kernel="wave", rng="sample"))
    _, rad_m, st_m = scene.render(cam, rtc.make_opts(kernel="mega", rng="sample"))
    img_o, img_p, img_w, img_m = rad_o / spp, rad_p / spp, rad_w / spp, rad_m / spp
    ok = np.isfinite(img_o).all(axis=2) & np.isfinite(img_w).all(axis=2) & np.isfinite(img_p).all(axis=2) & \
        np.isfinite(img_m).all(axis=2)
    assert ok.mean() > 0.99
    img_o, img_p, img_w, img_m = (np.where(ok[..., None], x, 0.0) for x in (img_o, img_p, img_w, img_m))
    floor = _rmse(img_o, img_p)
    assert (_rmse(img_w, img_o) <= 0.5 * floor + 1e-4).all(), (_rmse(img_w, img_o), floor)
    assert (_rmse(img_w, img_m) <= 0.25 * floor + 1e-4).all()
    bias = np.abs(np.clip(img_w, 0, 1).mean(axis=(0, 1)) - np.clip(img_o, 0, 1).mean(axis=(0, 1)))
    assert (bias <= 2.0 / 255.0).all(), bias
    assert st_w["samples"] == st_o["samples"]
    assert abs(st_w["segments"] - st_o["segments"]) <= 0.02 * st_o["segments"]
    assert st_w["launches"] > 8


def test_wavefront_is_deterministic_and_handles_pixel_streams(rtc, oracle, gpu_scene):
    sess, desc, scene = gpu_scene(1, 96, 32)
    cam = sess.camera
    opts = rtc.make_opts(kernel="wave", rng="sample")
    _, a, st_a = scene.render(cam, opts)
    _, b, st_b = scene.render(cam, opts)
    assert same_bits(a, b).all() and st_a["segments"] == st_b["segments"]
    _, lo, _ = scene.render(cam, rtc.make_opts(kernel="wave", rng="sample", sample_begin=0, sample_count=20))
    _, hi, _ = scene.render(cam, rtc.make_opts(kernel="wave", rng="sample", sample_begin=20, sample_count=12))
    assert np.allclose(lo + hi, a, rtol=2e-5, atol=1e-5)
    _, rad_px, _ = scene.render(cam, rtc.make_opts(kernel="wave", rng="pixel"))
    _, rad_o, _ = oracle.render(desc, cam, rtc.make_opts(kernel="exact", rng="pixel"), oracle.MATH_PORTABLE)
    _, rad_s, _ = oracle.render(desc, cam, rtc.make_opts(kernel="exact", rng="sample"), oracle.MATH_PORTABLE)
    assert (_rmse(rad_px / 32, rad_o / 32) <= 0.6 * _rmse(rad_s / 32, rad_o / 32) + 1e-4).all()


def test_pixel_rng_mode_on_mega(rtc, oracle, gpu_scene):
    """RT_RNG_PIXEL (the reference's own contract: one stream per pixel) also runs on the throughput kernel."""
    sess, desc, scene = gpu_scene(1, 64, 16)
    cam = sess.camera
    _, rad_o, _ = oracle.render(desc, cam, rtc.make_opts(kernel="exact", rng="pixel"), oracle.MATH_PORTABLE)
    _, rad_s, _ = oracle.render(desc, cam, rtc.make_opts(kernel="exact", rng="sample"), oracle.MATH_PORTABLE)
    _, rad_m, _ = scene.render(cam, rtc.make_opts(kernel="mega", rng="pixel"))
    assert (_rmse(rad_m / 16, rad_o / 16) <= 0.6 * _rmse(rad_s / 16, rad_o / 16) + 1e-4).all()


def test_render_is_deterministic(rtc, gpu_scene):
    """No float atomics: work items store partial sums, a second kernel adds them in chunk order."""
    sess, desc, scene = gpu_scene(7, 96, 64)
    opts = rtc.make_opts(kernel="mega", rng="sample")
    rgb_a, rad_a, st_a = scene.render(sess.camera, opts)
    rgb_b, rad_b, st_b = scene.render(sess.camera, opts)
    assert same_bits(rad_a, rad_b).all()
    assert np.array_equal(rgb_a, rgb_b)
    assert st_a["segments"] == st_b["segments"]


@pytest.mark.parametrize("scene_id", [1, 6, 7])
def test_all_megakernel_variants_agree_bit_for_bit(rtc, gpu_scene, scene_id, monkeypatch):
    """Free-running, lockstep-CTA and warp-ray-queue kernels, voted and plain node loops: the schedule of the work
    changes, the arithmetic of a sample and the order in which a work item sums its samples do not -- so the float
    radiance must be identical bit for bit, and so must the segment counts."""
    # (RT_BVH_PRE=0: a scene with an oversized primitive outside its tree can only be rendered by the kernels
    # instantiated for that, and every forced shape would silently fall back to them)
    monkeypatch.setenv("RT_BVH_PRE", "0")
    sess, desc, _ = gpu_scene(scene_id, 96, 32)
    scene = rtc.Scene(desc)
    monkeypatch.delenv("RT_BVH_PRE")
    opts = rtc.make_opts(kernel="mega", rng="sample")
    monkeypatch.setenv("RT_MEGA_CHUNKS", "2")  # same chunking for every variant
    monkeypatch.setenv("RT_MEGA_LEAN", "0")    # ... and the same instantiation (all features) of every kernel shape
    monkeypatch.setenv("RT_MEGA_NV", "0")
    monkeypatch.setenv("RT_MEGA_MINB", "8")
    rgb_ref, rad_ref, st_ref = scene.render(sess.camera, opts)
    variants = [("0", "6", "2"), ("0", "4", "0"), ("-1", "4", "2"), ("-1", "2", "2"), ("-1", "1", "4"),
                ("2", "6", "2"), ("4", "4", "0"),
                # RESUME kernels (a trip ends before its longest rays do; stragglers carry on in the next trip)
                # (4th: RT_MEGA_RESUME8; 5th: RT_MEGA_CTA8, the CTA-wide early end of a lockstep trip, 0 = off)
                ("-2", "2", "2", "1"), ("-2", "2", "2", "1", "0"), ("-2", "2", "2", "2", "6"), ("-2", "4", "2", "3", "1"),
                ("-3", "8", "2", "2"), ("-3", "8", "1", "6")]
    for nv, minb, thr8, *resume8 in variants:
        monkeypatch.setenv("RT_MEGA_NV", nv)
        monkeypatch.setenv("RT_MEGA_MINB", minb)
        monkeypatch.setenv("RT_MEGA_THR8", thr8)
        monkeypatch.setenv("RT_MEGA_RESUME8", resume8[0] if resume8 else "1")
        monkeypatch.setenv("RT_MEGA_CTA8", resume8[1] if len(resume8) > 1 else "3")
        rgb, rad, st = scene.render(sess.camera, opts)
        assert same_bits(rad, rad_ref).all(), (nv, minb, thr8, resume8)
        assert np.array_equal(rgb, rgb_ref)
        assert st["segments"] == st_ref["segments"] and st["samples"] == st_ref["samples"]
    # The LEAN instantiations (the same kernels without the code of features the scene lacks) against the full ones:
    # another instantiation is other machine code -- the compiler contracts other multiply-adds in the fast shading
    # arithmetic -- so a few paths fork on a last bit; the same samples otherwise.  Not bit-identical, but far inside
    # the noise: most pixels equal, no bias.
    for name in ("RT_MEGA_NV", "RT_MEGA_MINB", "RT_MEGA_THR8", "RT_MEGA_RESUME8", "RT_MEGA_CTA8"):
        monkeypatch.delenv(name, raising=False)
    rgb_f, rad_f, st_f = scene.render(sess.camera, opts)
    monkeypatch.delenv("RT_MEGA_LEAN")
    rgb_l, rad_l, st_l = scene.render(sess.camera, opts)
    assert st_l["samples"] == st_f["samples"] and abs(st_l["segments"] - st_f["segments"]) <= 0.005 * st_f["segments"]
    a, b = np.clip(rad_l / 32, 0, 1), np.clip(rad_f / 32, 0, 1)
    assert np.sqrt(((a - b) ** 2).mean()) < 0.03
    assert abs(a.mean() - b.mean()) < 2e-3
    scene.close()


def test_warp_item_pool_renders_the_same_bits_as_per_lane_draws(rtc, gpu_scene, monkeypatch):
    """The lean sphere kernel exists with the warp item pool (RT_MEGA_MINB=8: a warp draws the 32 items of a tile's
    chunk and hands them to its lanes) and with one draw per lane (RT_MEGA_MINB=10, 12): who renders which item
    changes, what an item sums does not -- identical radiance bits, counters and 8-bit frames; on a frame whose
    width is not a multiple of the tile (padding items) and with few and with many chunks."""
    monkeypatch.setenv("RT_BVH_PRE", "0")  # the kernels without the pre-walk primitives: FEATS = 0 in all three builds
    sess, desc, _ = gpu_scene(1, 101, 24)
    scene = rtc.Scene(desc)
    monkeypatch.delenv("RT_BVH_PRE")
    opts = rtc.make_opts(kernel="mega", rng="sample")
    monkeypatch.setenv("RT_MEGA_NV", "0")
    try:
        for chunks in ("1", "2", "0"):  # 0: the guided plan (many items of falling size)
            if chunks == "0":
                monkeypatch.delenv("RT_MEGA_CHUNKS")
            else:
                monkeypatch.setenv("RT_MEGA_CHUNKS", chunks)
            monkeypatch.setenv("RT_MEGA_MINB", "8")
            rgb_p, rad_p, st_p = scene.render(sess.camera, opts)
            assert st_p["samples"] == sess.width * sess.height * 24
            for minb in ("10", "12"):
                monkeypatch.setenv("RT_MEGA_MINB", minb)
                rgb, rad, st = scene.render(sess.camera, opts)
                assert same_bits(rad, rad_p).all(), (chunks, minb)
                assert np.array_equal(rgb, rgb_p)
                assert st["segments"] == st_p["segments"] and st["samples"] == st_p["samples"]
    finally:
        scene.close()


@pytest.mark.parametrize("scene_id", [0, 1, 4, 5, 6, 7])
def test_scene_build_options_agree_bit_for_bit(rtc, scene_factory, scene_id, monkeypatch):
    """Box leaves (six quads of a Box_new list in one leaf) against one leaf per quad, the one-quad-light fast path
    against the generic walk over world.lights, the 4-wide tree against the binary one, oversized primitives tested
    before the walk against the same primitives inside the tree (scenes 0, 1: ground spheres): other structure, other code,
    the same hits, the same draws -- the float radiance must be identical bit for bit."""
    sess, desc = scene_factory(scene_id, 96, 32)
    opts = rtc.make_opts(kernel="mega", rng="sample")
    results = []
    for env in ({}, {"RT_BVH_BOXES": "1"}, {"RT_ONE_LIGHT_OFF": "1"}, {"RT_BVH_WIDE": "1"},
                {"RT_BVH_WIDE": "1", "RT_BVH_BOXES": "1"}, {"RT_BVH_PRE": "0"}, {"RT_BVH_PRE": "0", "RT_BVH_WIDE": "1"}):
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        scene = rtc.Scene(desc)
        results.append(scene.render(sess.camera, opts))
        scene.close()
        for k in env:
            monkeypatch.delenv(k)
    for rgb, rad, st in results[1:]:
        assert same_bits(rad, results[0][1]).all()
        assert st["segments"] == results[0][2]["segments"]


def test_sample_ranges_add_up(rtc, gpu_scene):
    sess, desc, scene = gpu_scene(6, 64, 32)
    cam = sess.camera
    _, whole, st = scene.render(cam, rtc.make_opts(kernel="mega", rng="sample"))
    parts = np.zeros_like(whole)
    segs = 0
    for begin, count in ((0, 5), (5, 11), (16, 16)):
        _, part, st_p = scene.render(cam, rtc.make_opts(kernel="mega", rng="sample", sample_begin=begin,
                                                        sample_count=count))
        parts += part
        segs += st_p["segments"]
    assert segs == st["segments"]
    assert np.allclose(parts, whole, rtol=2e-5, atol=1e-5)


def test_full_size_book1_properties(rtc, gpu_scene):
    """BASELINE config 2 geometry (1200x675, depth 50) at 16 spp: size-independent properties."""
    sess, desc, scene = gpu_scene(1, 1200, 16)
    cam = sess.camera
    assert (cam.img_width, cam.img_height, cam.max_depth) == (1200, 675, 50)
    opts = rtc.make_opts(kernel="mega", rng="sample")
    rgb, rad, st = scene.render(cam, opts)
    assert st["samples"] == 1200 * 675 * 16
    assert 2.3 < st["segments"] / st["samples"] < 3.2  # 2.72 segments per sample
    assert st["nan_samples"] == 0 and np.isfinite(rad).all()
    # halves add up to the whole, and a second run is bit-identical
    _, lo, _ = scene.render(cam, rtc.make_opts(kernel="mega", rng="sample", sample_begin=0, sample_count=8))
    _, hi, _ = scene.render(cam, rtc.make_opts(kernel="mega", rng="sample", sample_begin=8, sample_count=8))
    assert np.allclose(lo + hi, rad, rtol=2e-5, atol=1e-5)
    _, again, _ = scene.render(cam, opts)
    assert same_bits(again, rad).all()
    # the exact kernel (bit-identical to the oracle) at the same size and seeds gives the same picture
    _, rad_e, st_e = scene.render(cam, rtc.make_opts(kernel="exact", rng="sample", sample_count=4))
    _, rad_4, _ = scene.render(cam, rtc.make_opts(kernel="mega", rng="sample", sample_count=4))
    assert (_rmse(rad_4 / 4, rad_e / 4) < 0.02).all()
    assert abs(np.clip(rad_4 / 4, 0, 1).mean() - np.clip(rad_e / 4, 0, 1).mean()) < 1.0 / 255


@pytest.mark.parametrize("scene_id,width,size,seg_lo,seg_hi", [(6, 600, (600, 600), 2.4, 3.5),
                                                               (7, 1000, (1000, 1000), 4.4, 6.3)])
def test_full_size_cornell_and_book2_properties(rtc, gpu_scene, scene_id, width, size, seg_lo, seg_hi):
    """BASELINE configs 3 and 4 geometry (600x600 Cornell with the mixture pdf, 1000x1000 Book 2 final; depth 50) at
    8 spp: the same size-independent properties as config 2 -- sample and segment accounting (2.93 and
    5.32 segments per sample), finite radiance, sample ranges add up, reruns are bit-identical, and the picture agrees
    with the exact kernel's (which is the oracle's, bit for bit)."""
    sess, desc, scene = gpu_scene(scene_id, width, 8)
    cam = sess.camera
    assert (cam.img_width, cam.img_height, cam.max_depth) == (size[0], size[1], 50)
    opts = rtc.make_opts(kernel="mega", rng="sample")
    rgb, rad, st = scene.render(cam, opts)
    assert st["samples"] == size[0] * size[1] * 8
    assert seg_lo < st["segments"] / st["samples"] < seg_hi
    assert np.isfinite(rad).all()
    if scene_id == 6:
        assert st["light_pdf_evals"] > 0  # the mixture pdf is exercised
    _, lo, _ = scene.render(cam, rtc.make_opts(kernel="mega", rng="sample", sample_begin=0, sample_count=3))
    _, hi, _ = scene.render(cam, rtc.make_opts(kernel="mega", rng="sample", sample_begin=3, sample_count=5))
    assert np.allclose(lo + hi, rad, rtol=2e-5, atol=1e-5)
    _, again, st2 = scene.render(cam, opts)
    assert same_bits(again, rad).all() and st2["segments"] == st["segments"]
    _, rad_e, _ = scene.render(cam, rtc.make_opts(kernel="exact", rng="sample", sample_count=2))
    _, rad_2, _ = scene.render(cam, rtc.make_opts(kernel="mega", rng="sample", sample_count=2))
    # Cornell at 2 spp is dominated by fireflies of the light sampling: compare clipped means and a loose RMSE
    assert abs(np.clip(rad_2 / 2, 0, 1).mean() - np.clip(rad_e / 2, 0, 1).mean()) < 2.0 / 255
    assert (_rmse(np.clip(rad_2 / 2, 0, 1), np.clip(rad_e / 2, 0, 1)) < 0.25).all()


def test_unknown_kernel_is_refused(rtc, gpu_scene):
    sess, desc, scene = gpu_scene(0)
    with pytest.raises(rtc.lib.RtError) as err:
        scene.render(sess.camera, rtc.make_opts(kernel=9))
    assert err.value.status in (rtc.abi.RT_ERR_UNSUPPORTED, rtc.abi.RT_ERR_INVALID)


@pytest.mark.parametrize("n_instances,with_quads", [(60, True), (200, False)])
def test_many_flattened_instances_hit_parity(rtc, scene_factory, n_instances, with_quads):
    """Dozens of Translate(RotateY(primitive)) chains with arbitrary angles and offsets (the reference's scenes have
    at most two): the rebuilt tree bounds every primitive in the world frame and tests it in its own -- hit distances
    and ids must still be the exact kernel's (= the reference's order of operations) bit for bit, from outside and
    from points inside the cloud."""
    from synth_scenes import instanced_scene

    _, base = scene_factory(1, 48, 4)
    desc = instanced_scene(rtc, base, n_instances, seed=11 + n_instances, with_quads=with_quads)
    scene = rtc.Scene(desc)
    try:
        rng = np.random.default_rng(5)
        n = 60000
        o = np.concatenate([np.tile(np.array([[0, 3, 40]], np.float32), (n // 2, 1)),
                            rng.uniform(-15, 15, (n - n // 2, 3)).astype(np.float32)])
        target = rng.uniform(-16, 16, (n, 3)).astype(np.float32)
        d = (target - o).astype(np.float32)
        ref = scene.hit_batch(o, d, 1e-3, np.inf, kernel="exact")
        got = scene.hit_batch(o, d, 1e-3, np.inf, kernel="mega")
        assert np.isfinite(ref["t"]).mean() > 0.2
        assert same_bits(got["t"], ref["t"]).all()
        assert (got["node"] == ref["node"]).all()
        hit = np.isfinite(ref["t"])
        assert same_bits(got["p"][hit], ref["p"][hit]).all() and same_bits(got["normal"][hit], ref["normal"][hit]).all()
    finally:
        scene.close()
