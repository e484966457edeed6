"""Edge cases of the render path through the C ABI, every kernel against the oracle: empty world, single-primitive
world, image sizes that do not fill the 8x4 pixel tiles, zero depth, a single sample, empty ray batches, sample ranges
at the ends.  (The reference has no tests of its own; these are the degenerate inputs its loops admit:
raytracing.c:40-41 depth <= 0, :100 spp loop, hittable.c:62-69 empty list.)"""
from __future__ import annotations

import numpy as np
import pytest

from conftest import same_bits

pytestmark = pytest.mark.gpu


def _clone(desc, rtc, mutate):
    clone = rtc.abi.SceneDesc.from_bytes(desc.to_bytes())
    mutate(clone)
    return clone


def _root_list(d, rtc):
    root = d.c.objects_root
    assert d.nodes["kind"][root] == rtc.abi.NODE_LIST
    return root


def test_empty_world_renders_the_background(rtc, oracle, scene_factory):
    """HittableList_hit over zero objects (hittable.c:62-69): every path ends on the background after one segment."""
    sess, desc = scene_factory(2, 40, 3)

    def empty(d):
        d.nodes["b"][_root_list(d, rtc)] = 0

    bare = _clone(desc, rtc, empty)
    cam = sess.camera
    _, rad_o, st_o = oracle.render(bare, cam, rtc.make_opts(kernel="exact", rng="pixel"), oracle.MATH_PORTABLE)
    scene = rtc.Scene(bare)
    try:
        _, rad_e, st_e = scene.render(cam, rtc.make_opts(kernel="exact", rng="pixel"))
        assert same_bits(rad_e, rad_o).all() and st_e["segments"] == st_o["segments"]
        for kernel in ("mega", "wave"):
            _, rad, st = scene.render(cam, rtc.make_opts(kernel=kernel, rng="sample"))
            assert st["segments"] == st["samples"] == cam.img_width * cam.img_height * 3
            assert np.allclose(rad / 3, np.asarray(cam.background, np.float32)[None, None, :], rtol=1e-6), kernel
        n = 1000
        o = np.zeros((n, 3), np.float32)
        d = np.random.default_rng(1).normal(size=(n, 3)).astype(np.float32)
        for kernel in ("exact", "mega"):
            hits = scene.hit_batch(o, d, 1e-3, np.inf, kernel=kernel)
            assert np.isinf(hits["t"]).all() and (hits["node"] == rtc.abi.RT_NONE).all()
    finally:
        scene.close()


def test_single_primitive_world(rtc, oracle, scene_factory):
    """One sphere: the rebuilt structure is a root that is a leaf (no inner node at all)."""
    sess, desc = scene_factory(2, 40, 3)

    def one(d):
        d.nodes["b"][_root_list(d, rtc)] = 1

    single = _clone(desc, rtc, one)
    info = rtc.inspect(single)
    assert info["bvh_nodes"] == 0 and info["spheres"] == 1
    cam = sess.camera
    scene = rtc.Scene(single)
    try:
        _, rad_o, _ = oracle.render(single, cam, rtc.make_opts(kernel="exact", rng="pixel"), oracle.MATH_PORTABLE)
        _, rad_e, _ = scene.render(cam, rtc.make_opts(kernel="exact", rng="pixel"))
        assert same_bits(rad_e, rad_o).all()
        rng = np.random.default_rng(2)
        o = np.tile(np.asarray(cam.origin, np.float32), (4000, 1))
        d = rng.normal(size=(4000, 3)).astype(np.float32)
        ref = scene.hit_batch(o, d, 1e-3, np.inf, kernel="exact")
        got = scene.hit_batch(o, d, 1e-3, np.inf, kernel="mega")
        assert same_bits(got["t"], ref["t"]).all() and (got["node"] == ref["node"]).all()
        assert np.isfinite(ref["t"]).any()
    finally:
        scene.close()


@pytest.mark.parametrize("width,aspect", [(1, 1.0), (7, 7.0 / 3.0), (33, 33.0 / 5.0), (50, 50.0 / 27.0)])
def test_ragged_image_sizes(rtc, scene_factory, width, aspect):
    """Widths and heights that leave the last 8x4 pixel tile column / row partly outside the image."""
    sess, desc = scene_factory(1, width, 4, aspect=aspect)
    cam = sess.camera
    scene = rtc.Scene(desc)
    try:
        _, rad_e, st_e = scene.render(cam, rtc.make_opts(kernel="exact", rng="sample"))
        for kernel in ("mega", "wave"):
            _, rad, st = scene.render(cam, rtc.make_opts(kernel=kernel, rng="sample"))
            assert rad.shape == (cam.img_height, cam.img_width, 3)
            assert st["samples"] == cam.img_width * cam.img_height * 4 == st_e["samples"]
            assert np.isfinite(rad).all() and (rad >= 0).all()
            # every pixel received its samples: the sky gradient is never black
            assert (rad.sum(axis=2) > 0).all(), kernel
    finally:
        scene.close()


def test_zero_depth_and_single_sample(rtc, scene_factory):
    """max_depth = 0: Camera_ray_color returns black at once (raytracing.c:40-41), no segment is traced."""
    sess, desc = scene_factory(1, 48, 5)
    cam = sess.camera
    cam.max_depth = 0  # (the session's own `depth=0` means "the scene's default")
    scene = rtc.Scene(desc)
    try:
        for kernel in ("exact", "mega", "wave"):
            _, rad, st = scene.render(cam, rtc.make_opts(kernel=kernel, rng="sample"))
            assert st["segments"] == 0 and st["samples"] == cam.img_width * cam.img_height * 5, kernel
            assert not rad.any()
    finally:
        scene.close()
    sess1, desc1 = scene_factory(1, 48, 1)
    scene = rtc.Scene(desc1)
    try:
        _, rad_m, st_m = scene.render(sess1.camera, rtc.make_opts(kernel="mega", rng="sample"))
        _, rad_e, st_e = scene.render(sess1.camera, rtc.make_opts(kernel="exact", rng="sample"))
        assert st_m["samples"] == st_e["samples"] == sess1.camera.img_width * sess1.camera.img_height
        assert np.isfinite(rad_m).all()
    finally:
        scene.close()


def test_sample_range_ends_and_empty_batches(rtc, scene_factory):
    sess, desc = scene_factory(6, 32, 8)
    cam = sess.camera
    scene = rtc.Scene(desc)
    try:
        _, whole, _ = scene.render(cam, rtc.make_opts(kernel="mega", rng="sample"))
        _, first, st_f = scene.render(cam, rtc.make_opts(kernel="mega", rng="sample", sample_begin=0, sample_count=1))
        _, rest, st_r = scene.render(cam, rtc.make_opts(kernel="mega", rng="sample", sample_begin=1, sample_count=7))
        assert st_f["samples"] == 32 * 32 and st_r["samples"] == 32 * 32 * 7
        assert np.allclose(first + rest, whole, rtol=2e-5, atol=1e-5)
        # explicit ranges are not tied to the camera's sample count (progressive accumulation): [4, 12) is legal and
        # its first half is the camera's second half; a negative start is refused
        _, beyond, st_b = scene.render(cam, rtc.make_opts(kernel="mega", rng="sample", sample_begin=4, sample_count=8))
        assert st_b["samples"] == 32 * 32 * 8 and np.isfinite(beyond).all()
        with pytest.raises(Exception):
            scene.render(cam, rtc.make_opts(kernel="mega", rng="sample", sample_begin=-1, sample_count=2))
        empty = np.zeros((0, 3), np.float32)
        for kernel in ("exact", "mega"):
            hits = scene.hit_batch(empty, empty, 1e-3, np.inf, kernel=kernel)
            assert hits["t"].shape == (0,)
    finally:
        scene.close()
