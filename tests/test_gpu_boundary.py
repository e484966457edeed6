"""The drop-in boundary on the GPU: Camera_render (World + Camera in host memory -> 8-bit image in host memory), the
reference's command line, TIFF output, and the multi-GPU entry points."""
from __future__ import annotations

import json
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


def test_camera_render_over_two_gpus_in_one_process(rtc):
    """The literal drop-in at N > 1: `RT_GPUS=2 rtc_main ...` = Camera_render -> rt_render_multi -> one in-process
    ncclReduce (communicator from ncclCommInitAll, kept between calls); with RT_MULTI_NCCL=0 the frames are gathered
    over peer copies instead.  Both must give the one-GPU image (same per-(pixel, sample) streams; the float sums are
    added in another order: at most one 8-bit level)."""
    if rtc.lib.b200().rt_device_count() < 2:
        pytest.skip("needs 2 GPUs")
    exe = rtc.lib.BUILD_DIR / "rtc_main"
    images = {}
    with tempfile.TemporaryDirectory() as tmp:
        for name, env in (("one", {"RT_GPUS": "1"}), ("nccl", {"RT_GPUS": "2"}),
                          ("peer", {"RT_GPUS": "2", "RT_MULTI_NCCL": "0"})):
            proc = subprocess.run([str(exe), "7", "--width", "96", "--spp", "16", "--out", f"{name}.tiff"], cwd=tmp,
                                  capture_output=True, text=True, env=dict(os.environ, RT_STATS="1", **env))
            assert proc.returncode == 0, proc.stderr
            stats = json.loads([l for l in proc.stderr.splitlines() if l.startswith("{")][-1])
            assert stats["n_gpus"] == int(env["RT_GPUS"]) and stats["samples"] == 96 * 96 * 16
            images[name] = _tiff_payload(Path(tmp) / f"{name}.tiff", 96, 96)
    for name in ("nccl", "peer"):
        assert np.abs(images[name].astype(int) - images["one"].astype(int)).max() <= 1
    assert np.array_equal(images["nccl"], images["peer"]) or (images["nccl"] != images["peer"]).mean() < 0.01


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


def _tiff_payload(path, width, height):
    data = Path(path).read_bytes()
    assert data[:4] == b"II*\x00"
    assert int.from_bytes(data[8 + 2 + 8:8 + 2 + 10], "little") == width      # ImageWidth, first IFD entry
    assert int.from_bytes(data[8 + 2 + 12 + 8:8 + 2 + 12 + 10], "little") == height
    return np.frombuffer(data[-width * height * 3:], dtype=np.uint8).reshape(height, width, 3)


def test_cli_named_options(rtc):
    """rtc_main's options beyond the reference's positional arguments (src/main.c:280-292): --width / --spp (the
    settings without the argv quirk), --depth, --aspect, --lights-prob, --out; the frame must be the one a Session
    with the same settings renders, byte for byte."""
    exe = rtc.lib.BUILD_DIR / "rtc_main"
    with tempfile.TemporaryDirectory() as tmp:
        env = dict(os.environ, RT_STATS="1")
        proc = subprocess.run([str(exe), "6", "--width", "64", "--spp", "8", "--depth", "5", "--aspect", "2.0",
                               "--lights-prob", "0.3", "--out", "cornell.tiff"], cwd=tmp, capture_output=True,
                              text=True, env=env)
        assert proc.returncode == 0, proc.stderr
        assert not (Path(tmp) / "output.tiff").exists()
        got = _tiff_payload(Path(tmp) / "cornell.tiff", 64, 32)
        stats = json.loads([l for l in proc.stderr.splitlines() if l.startswith("{")][-1])
    assert stats["samples"] == 64 * 32 * 8
    sess = rtc.Session(6, width=64, spp=8, depth=5, aspect=2.0, lights_prob=0.3)
    want, _, st = sess.camera_render(rtc.make_opts(kernel="mega", rng="sample"))
    assert (sess.width, sess.height) == (64, 32)
    assert np.array_equal(got, want)
    assert stats["segments"] == st["segments"]
    # depth is honoured: no path is longer than 5 segments
    assert st["segments"] <= 5 * st["samples"]
    # and the positional quirk is still the reference's: `6 64 8` (three arguments) sets the width, not the spp
    with tempfile.TemporaryDirectory() as tmp:
        proc = subprocess.run([str(exe), "6", "32", "8", "--depth", "2"], cwd=tmp, capture_output=True, text=True,
                              env=dict(os.environ, RT_STATS="1"))
        assert proc.returncode == 0, proc.stderr
        stats = json.loads([l for l in proc.stderr.splitlines() if l.startswith("{")][-1])
        assert stats["samples"] == 32 * 32 * 100


def test_cli_progressive_passes_and_checkpoint_resume(rtc):
    """--passes K renders the samples in K sample-range passes and --checkpoint saves the float sum after each; a
    run killed after some passes and resumed must give the image of the uninterrupted run (the per-(pixel, sample)
    streams make sample ranges additive)."""
    exe = rtc.lib.BUILD_DIR / "rtc_main"
    base = ["1", "--width", "96", "--spp", "24", "--depth", "20"]
    with tempfile.TemporaryDirectory() as tmp:
        one = subprocess.run([str(exe)] + base + ["--out", "one.tiff"], cwd=tmp, capture_output=True, text=True)
        assert one.returncode == 0, one.stderr
        prog = subprocess.run([str(exe)] + base + ["--passes", "4", "--checkpoint", "ck.bin", "--out", "prog.tiff"],
                              cwd=tmp, capture_output=True, text=True)
        assert prog.returncode == 0, prog.stderr
        a, b = _tiff_payload(Path(tmp) / "one.tiff", 96, 54), _tiff_payload(Path(tmp) / "prog.tiff", 96, 54)
        # same samples, summed in another order: at most the last bit of a sum, i.e. at most one 8-bit level
        assert np.abs(a.astype(int) - b.astype(int)).max() <= 1
        assert (a != b).mean() < 0.01
        # a run that stops half way (--stop-after 2: two of the four passes of 6 samples), then is resumed
        half = subprocess.run([str(exe)] + base + ["--passes", "4", "--stop-after", "2", "--checkpoint", "half.bin",
                                                    "--out", "half.tiff"], cwd=tmp, capture_output=True, text=True)
        assert half.returncode == 0, half.stderr
        hdr = np.fromfile(Path(tmp) / "half.bin", dtype="<u4", count=5)
        assert hdr[0] == 0x54435452 and tuple(hdr[1:4]) == (96, 54, 24) and hdr[4] == 12
        res = subprocess.run([str(exe)] + base + ["--passes", "4", "--checkpoint", "half.bin", "--resume", "--out",
                                                   "res.tiff"], cwd=tmp, capture_output=True, text=True)
        assert res.returncode == 0 and "Resuming at sample 12 of 24" in res.stderr, res.stderr
        c = _tiff_payload(Path(tmp) / "res.tiff", 96, 54)
        assert np.array_equal(c, b)


def test_camera_render_falls_back_when_the_rebuilt_structure_cannot_hold_the_scene(rtc):
    """300 distinct Translate(RotateY(...)) chains exceed the 255 the throughput kernels can tag; the reference has no
    such limit.  Camera_render must say so and render with the reference-order kernel instead of aborting."""
    from conftest import ROOT

    build = rtc.lib.BUILD_DIR
    with tempfile.TemporaryDirectory() as tmp:
        exe = Path(tmp) / "many_chains"
        subprocess.run(["gcc", "-std=gnu11", "-O1", f"-I{ROOT / 'include'}", str(ROOT / "tests" / "c" / "many_chains.c"),
                        "-o", str(exe), f"-L{build}", "-lrtc_host", "-lrtc_b200", f"-Wl,-rpath,{build}", "-lm"],
                       check=True)
        few = subprocess.run([str(exe), "40"], capture_output=True, text=True)
        assert few.returncode == 0 and "RT_KERNEL=exact" not in few.stderr, few.stderr
        many = subprocess.run([str(exe), "300"], capture_output=True, text=True)
        assert many.returncode == 0, many.stderr
        assert "RT_KERNEL=exact" in many.stderr and "255" in many.stderr
        w, h, total = (int(v) for v in many.stdout.split())
        assert (w, h) == (64, 32) and total > 64 * 32 * 3 * 20  # a lit frame, not a black one


def test_camera_render_reuses_the_resident_scene(rtc):
    """Camera_render goes through rt_scene_acquire / rt_scene_release: the second call on the same World must not
    rebuild and upload (same pixels either way; the scene cache is keyed by the description's content)."""
    import time

    sess = rtc.Session(7, width=64, spp=2)
    opts = rtc.make_opts(kernel="mega", rng="sample")
    rgb_a, _, _ = sess.camera_render(opts)
    t0 = time.perf_counter()
    rgb_b, _, _ = sess.camera_render(opts)
    warm = time.perf_counter() - t0
    assert np.array_equal(rgb_a, rgb_b)
    rtc.lib.b200().rt_shutdown()  # drops the idle cached scenes
    t0 = time.perf_counter()
    rgb_c, _, _ = sess.camera_render(opts)
    cold = time.perf_counter() - t0
    assert np.array_equal(rgb_a, rgb_c)
    assert warm < cold  # the 2.5 MB scene with its 3 400 primitives is built and uploaded only when it is not resident
    other = rtc.Session(7, width=64, spp=2, lights_prob=0.25)  # same World content, another camera: still a hit
    rgb_d, _, _ = other.camera_render(opts)
    assert rgb_d.shape == rgb_a.shape
