"""The drop-in boundary on the GPU: Camera_render (World + Camera in host memory -> 8-bit image in host memory), the
reference's command line, TIFF output, and the multi-GPU entry points."""
from __future__ import annotations

import os
import subprocess
import tempfile
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_camera_render_equals_the_abi_path(rtc, scene_factory):
    """Session.camera_render = flatten + rt_scene_create + rt_render + destroy (what Camera_render does) gives the
    same bytes as a resident rt_scene."""
    sess, desc = scene_factory(6, 64, 16)
    opts = rtc.make_opts(kernel="mega", rng="sample")
    rgb_a, rad_a, st_a = sess.camera_render(opts, want_radiance=True)
    scene = rtc.Scene(desc)
    rgb_b, rad_b, st_b = scene.render(sess.camera, opts)
    scene.close()
    assert np.array_equal(rgb_a, rgb_b)
    assert np.array_equal(rad_a.view(np.uint32), rad_b.view(np.uint32))
    assert st_a["segments"] == st_b["segments"] > 0


def test_literal_camera_render_call(rtc, scene_factory, oracle):
    """void Camera_render(const Camera*, const World*, uint8_t*) -- include/raytracing.h:41 -- with RT_KERNEL=exact and
    RT_RNG=pixel in the environment reproduces the oracle's image byte for byte."""
    sess, desc = scene_factory(0, 48, 4)
    rgb = np.zeros((sess.height, sess.width, 3), np.uint8)
    old = {k: os.environ.get(k) for k in ("RT_KERNEL", "RT_RNG", "RT_GPUS")}
    os.environ.update(RT_KERNEL="exact", RT_RNG="pixel")
    os.environ.pop("RT_GPUS", None)
    try:
        rtc.lib.host().rtc_session_camera_render(sess._p, rgb.ctypes.data)
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    want, _, _ = oracle.render(desc, sess.camera, rtc.make_opts(kernel="exact", rng="pixel"), oracle.MATH_PORTABLE)
    assert np.array_equal(rgb, want)


def test_cli_writes_the_reference_tiff(rtc, scene_factory):
    """`rtc_main 0 40 3 x` == `./main 0 40 3 x` of the reference: 40 px wide, 3 spp, output.tiff in the cwd whose
    payload (last W*H*3 bytes, src/tiff.c:47) is the rendered image."""
    exe = rtc.lib.BUILD_DIR / "rtc_main"
    with tempfile.TemporaryDirectory() as tmp:
        env = dict(os.environ, RT_KERNEL="exact", RT_RNG="pixel")
        proc = subprocess.run([str(exe), "0", "40", "3", "x"], cwd=tmp, capture_output=True, text=True, env=env)
        assert proc.returncode == 0, proc.stderr
        assert "Done" in proc.stderr and "Took" in proc.stderr
        data = (Path(tmp) / "output.tiff").read_bytes()
    sess, desc = scene_factory(0, 40, 3)
    assert data[:4] == b"II*\x00"
    payload = np.frombuffer(data[-sess.width * sess.height * 3:], dtype=np.uint8).reshape(sess.height, sess.width, 3)
    scene = rtc.Scene(desc)
    rgb, _, _ = scene.render(sess.camera, rtc.make_opts(kernel="exact", rng="pixel"))
    scene.close()
    assert np.array_equal(payload, rgb)
    # the width-only form (3 positional arguments) must leave spp at the default 100: just check it parses
    with tempfile.TemporaryDirectory() as tmp:
        proc = subprocess.run([str(exe), "2", "16", "1"], cwd=tmp, capture_output=True, text=True)
        assert proc.returncode == 0, proc.stderr


def test_reference_main_c_runs_on_the_backend_unchanged(rtc, scene_factory):
    """oracle/_ref/main_dropin = the reference's OWN src/main.c (scene builders, CLI, timing, TIFF call), compiled
    unchanged against include/*.h and linked to librtc_host.so + librtc_b200.so (option A).  Its
    output.tiff must carry the image this backend renders for the same scene."""
    from conftest import ROOT

    exe = ROOT / "oracle" / "_ref" / "main_dropin"
    if not exe.exists():
        pytest.skip("oracle/_ref/main_dropin not built (needs /root/reference at build time)")
    with tempfile.TemporaryDirectory() as tmp:
        env = dict(os.environ, RT_KERNEL="exact", RT_RNG="pixel")
        proc = subprocess.run([str(exe), "6", "40", "2", "x"], cwd=tmp, capture_output=True, text=True, env=env)
        assert proc.returncode == 0, proc.stderr
        data = (Path(tmp) / "output.tiff").read_bytes()
    sess, desc = scene_factory(6, 40, 2)
    payload = np.frombuffer(data[-sess.width * sess.height * 3:], dtype=np.uint8).reshape(sess.height, sess.width, 3)
    scene = rtc.Scene(desc)
    rgb, _, _ = scene.render(sess.camera, rtc.make_opts(kernel="exact", rng="pixel"))
    scene.close()
    assert np.array_equal(payload, rgb)


def test_render_multi_one_gpu_equals_render(rtc, scene_factory):
    sess, desc = scene_factory(5, 64, 16)
    opts = rtc.make_opts(kernel="mega", rng="sample")
    rgb_m, rad_m, st_m = rtc.render_multi(desc, sess.camera, opts, 1)
    scene = rtc.Scene(desc)
    rgb_s, rad_s, st_s = scene.render(sess.camera, opts)
    scene.close()
    assert np.array_equal(rgb_m, rgb_s)
    assert np.array_equal(rad_m.view(np.uint32), rad_s.view(np.uint32))
    assert st_m["segments"] == st_s["segments"]


def test_render_multi_shards_by_sample_range(rtc, scene_factory):
    n = rtc.lib.b200().rt_device_count()
    if n < 2:
        pytest.skip("needs 2 GPUs")
    sess, desc = scene_factory(6, 64, 32)
    opts = rtc.make_opts(kernel="mega", rng="sample")
    rgb_2, rad_2, st_2 = rtc.render_multi(desc, sess.camera, opts, 2)
    rgb_1, rad_1, st_1 = rtc.render_multi(desc, sess.camera, opts, 1)
    assert st_2["segments"] == st_1["segments"]
    assert np.allclose(rad_2, rad_1, rtol=2e-5, atol=1e-5)
    assert (np.abs(rgb_2.astype(int) - rgb_1.astype(int)) <= 1).all()


def test_dist_renderer_single_rank(rtc, scene_factory):
    """The one-process-per-GPU driver with a world of one: rt_render_device on torch's stream + tonemap + D2H."""
    import importlib

    dist = importlib.import_module("ray-tracing-c_b200.dist")
    sess, desc = scene_factory(1, 64, 8)
    r = dist.DistRenderer(desc, sess.camera, device=0)
    rgb, st = r.render(kernel="mega")
    scene = rtc.Scene(desc)
    want, _, st_w = scene.render(sess.camera, rtc.make_opts(kernel="mega", rng="sample"))
    scene.close()
    r.close()
    assert np.array_equal(rgb, want)
    assert st["segments"] == st_w["segments"]


def test_fp32_probe_reports_a_plausible_peak(rtc):
    tf = rtc.fp32_peak_tflops()
    assert 20.0 < tf < 120.0
