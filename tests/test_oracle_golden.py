"""Pins the oracle (oracle/rtc_oracle.c) and the host layer (constructors, BVH build, scene builders, Camera_init,
flattener) against what the UNMODIFIED reference produced: the committed fixtures in tests/golden/ (made by
tests/golden/make_golden.py from oracle/_ref/ref_tool) and, where /root/reference exists, the live tool.

Bar: byte-identical images, bit-identical hit records, identical flattened blobs."""
from __future__ import annotations

import hashlib
import subprocess
import tempfile
from pathlib import Path

import numpy as np
import pytest

from conftest import same_bits

SCENES = list(range(8))


@pytest.mark.parametrize("scene_id", SCENES)
def test_host_builders_match_reference_graph(rtc, golden, scene_factory, scene_id):
    """Our constructors + scene builders + BVH build produce, through the same flattener, the very blob the
    reference's own main.c / hittable.c produce (draw-for-draw identical scene, identical tree)."""
    g = golden(scene_id)
    sess, desc = scene_factory(scene_id, int(g["width"]), int(g["spp"]))
    digest = hashlib.sha256(desc.to_bytes()).digest()
    assert digest == bytes(g["blob_sha256"].tobytes())


@pytest.mark.parametrize("scene_id", SCENES)
def test_camera_init_matches_reference(rtc, golden, scene_factory, scene_id):
    g = golden(scene_id)
    sess, _ = scene_factory(scene_id, int(g["width"]), int(g["spp"]))
    assert sess.camera_raw() == bytes(g["camera_raw"].tobytes())


@pytest.mark.parametrize("scene_id", SCENES)
def test_oracle_render_is_byte_identical_to_reference(rtc, oracle, golden, scene_factory, scene_id):
    g = golden(scene_id)
    sess, desc = scene_factory(scene_id, int(g["width"]), int(g["spp"]))
    rgb8, _, stats = oracle.render(desc, sess.camera, rtc.make_opts(kernel="exact", rng="pixel"), oracle.MATH_LIBM)
    assert rgb8.shape == g["rgb8"].shape
    assert np.array_equal(rgb8, g["rgb8"])
    assert stats["segments"] > 0


@pytest.mark.parametrize("scene_id", SCENES)
def test_oracle_hits_are_bit_identical_to_reference(rtc, oracle, golden, scene_factory, scene_id):
    g = golden(scene_id)
    _, desc = scene_factory(scene_id, int(g["width"]), int(g["spp"]))
    out = oracle.hit_batch(desc, g["rays_o"], g["rays_d"], rng=g["rays_rng"], math=oracle.MATH_LIBM)
    hit = g["hit_hit"].astype(bool)
    assert np.array_equal(out["node"] != rtc.abi.RT_NONE, hit)
    assert same_bits(out["t"][hit], g["hit_t"][hit]).all()
    assert same_bits(out["p"][hit], g["hit_p"][hit]).all()
    assert same_bits(out["normal"][hit], g["hit_normal"][hit]).all()
    assert same_bits(out["uv"][hit, 0], g["hit_u"][hit]).all()
    assert same_bits(out["uv"][hit, 1], g["hit_v"][hit]).all()
    assert np.array_equal(out["front"][hit], g["hit_front"][hit].astype(np.uint8))
    assert np.array_equal(out["material"][hit], g["hit_material"][hit])
    assert np.array_equal(out["rng"][:, 0], g["hit_state"])  # every draw of a ConstantMedium, hit or miss


@pytest.mark.parametrize("scene_id", SCENES)
def test_portable_math_changes_nothing_visible(rtc, oracle, golden, scene_factory, scene_id):
    """PORTABLE mode (what the GPU's exact kernel reproduces bit for bit) vs LIBM mode (== the reference): same hits
    bit for bit except where a libm result feeds the record (sphere u,v; medium distance)."""
    g = golden(scene_id)
    _, desc = scene_factory(scene_id, int(g["width"]), int(g["spp"]))
    a = oracle.hit_batch(desc, g["rays_o"], g["rays_d"], rng=g["rays_rng"], math=oracle.MATH_LIBM)
    b = oracle.hit_batch(desc, g["rays_o"], g["rays_d"], rng=g["rays_rng"], math=oracle.MATH_PORTABLE)
    n_media = int((desc.nodes["kind"] == rtc.abi.NODE_MEDIUM).sum())
    if n_media == 0:
        assert np.array_equal(a["node"], b["node"])
        assert same_bits(a["t"], b["t"]).all()
    else:
        agree = a["node"] == b["node"]
        assert agree.mean() > 0.995
        ok = agree & np.isfinite(a["t"])
        assert np.allclose(a["t"][ok], b["t"][ok], rtol=2e-6)
    hit = (a["node"] != rtc.abi.RT_NONE) & (a["node"] == b["node"])
    assert np.nanmax(np.abs(a["uv"][hit] - b["uv"][hit]), initial=0.0) < 1e-6


@pytest.mark.parametrize("scene_id", [0, 1, 5, 6, 7])
def test_live_reference_matches_golden(rtc, ref_tool, golden, scene_id):
    """Where the reference sources exist: the committed fixture is what the reference produces today."""
    g = golden(scene_id)
    with tempfile.TemporaryDirectory() as tmp:
        subprocess.run([str(ref_tool), "render", str(scene_id), str(int(g["width"])), str(int(g["spp"])),
                        str(int(g["depth"])), "0", "out.rgb8", "camera.bin"], cwd=tmp, check=True,
                       stderr=subprocess.DEVNULL)
        rgb8 = np.fromfile(Path(tmp) / "out.rgb8", dtype=np.uint8).reshape(g["rgb8"].shape)
    assert np.array_equal(rgb8, g["rgb8"])


def test_pcg32_known_answers(oracle):
    """pcg-c-basic demo vector for seed (42, 54) -- /root/reference/src/pcg32.c:3-17."""
    want = [0xA15C02B7, 0x7B47F409, 0xBA1D3330, 0x83D2F293, 0xBFA4784B, 0xCBED606E]
    assert [int(v) for v in oracle.pcg32(42, 54, 6)] == want


def test_pcg32_advance_equals_stepping(rtc, oracle):
    """pcg32_advance (host extension, O(log n) jump-ahead over /root/reference/src/pcg32.c:11-17's single step): after
    advancing by n the stream continues with the oracle's (n+1)-th value -- for small, large and power-of-two n."""
    import ctypes as C

    class PCG32(C.Structure):
        _fields_ = [("state", C.c_uint64), ("inc", C.c_uint64)]

    host = rtc.lib.host()
    host.pcg32_seed.argtypes = [C.POINTER(PCG32), C.c_uint64, C.c_uint64]
    host.pcg32_advance.argtypes = [C.POINTER(PCG32), C.c_uint64]
    host.pcg32_u32.argtypes = [C.POINTER(PCG32)]
    host.pcg32_u32.restype = C.c_uint32
    want = [int(v) for v in oracle.pcg32(42, 54, 5000)]
    for n in (0, 1, 2, 3, 6, 64, 1000, 4095, 4096):
        g = PCG32()
        host.pcg32_seed(C.byref(g), 42, 54)
        host.pcg32_advance(C.byref(g), n)
        assert [host.pcg32_u32(C.byref(g)) for _ in range(4)] == want[n:n + 4], n
    # the affine map composes: advance(a) then advance(b) == advance(a + b), and 2^64 draws are the whole period
    a, b = PCG32(), PCG32()
    host.pcg32_seed(C.byref(a), 7, 9), host.pcg32_seed(C.byref(b), 7, 9)
    host.pcg32_advance(C.byref(a), 123456789), host.pcg32_advance(C.byref(a), 2 ** 40 + 17)
    host.pcg32_advance(C.byref(b), 123456789 + 2 ** 40 + 17)
    assert a.state == b.state
    host.pcg32_advance(C.byref(b), 2 ** 64 - 1), host.pcg32_u32(C.byref(b))
    assert a.state == b.state


def test_oracle_sample_mode_equals_pixel_mode_for_first_sample(rtc, oracle, scene_factory):
    """RT_RNG_SAMPLE seeds ((s<<32)|(17+j), 23+i): sample 0 is the reference's own stream."""
    sess, desc = scene_factory(0, 48, 1)
    cam = sess.camera
    _, a, _ = oracle.render(desc, cam, rtc.make_opts(kernel="exact", rng="pixel"), oracle.MATH_LIBM)
    _, b, _ = oracle.render(desc, cam, rtc.make_opts(kernel="exact", rng="sample"), oracle.MATH_LIBM)
    assert same_bits(a, b).all()


def test_oracle_sample_ranges_add_up(rtc, oracle, scene_factory):
    """Per-sample streams make a pixel's samples independent of how the range is split (multi-GPU sharding)."""
    sess, desc = scene_factory(6, 32, 8)
    cam = sess.camera
    _, whole, _ = oracle.render(desc, cam, rtc.make_opts(kernel="exact", rng="sample"), oracle.MATH_PORTABLE)
    _, lo, _ = oracle.render(desc, cam, rtc.make_opts(kernel="exact", rng="sample", sample_begin=0, sample_count=3),
                             oracle.MATH_PORTABLE)
    _, hi, _ = oracle.render(desc, cam, rtc.make_opts(kernel="exact", rng="sample", sample_begin=3, sample_count=5),
                             oracle.MATH_PORTABLE)
    assert np.allclose(lo + hi, whole, rtol=1e-5, atol=1e-6)
