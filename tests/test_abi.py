"""The C-ABI boundary without a GPU: the libraries load, export every declared symbol, the ctypes mirror matches the
header's layouts, malformed scenes are rejected, and compute entry points refuse to run without a device (there is
no CPU fallback to fall into)."""
from __future__ import annotations

import ctypes as C
import re
import subprocess
import tempfile
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent


def test_libraries_export_every_declared_symbol(rtc):
    header = (ROOT / "include" / "rtc_b200.h").read_text()
    declared = set(re.findall(r"\b(rt_[a-z0-9_]+)\s*\(", header))
    declared -= {"rt_status", "rt_scene_desc"}  # words followed by "(" in comments
    assert declared == set(rtc.lib.B200_SYMBOLS), declared ^ set(rtc.lib.B200_SYMBOLS)
    b200 = rtc.lib.b200()
    for name in rtc.lib.B200_SYMBOLS:
        assert hasattr(b200, name), name
    host = rtc.lib.host()
    for name in rtc.lib.HOST_SYMBOLS:
        assert hasattr(host, name), name
    assert b200.rt_abi_version() == rtc.abi.RT_ABI_VERSION


def test_ctypes_mirror_matches_header_layout(rtc):
    """sizeof/offsetof through a tiny C program compiled against include/rtc_b200.h."""
    prog = r"""
    #include "rtc_b200.h"
    #include <stdio.h>
    #include <stddef.h>
    int main(void) {
      printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu %zu\n", sizeof(rt_node), sizeof(rt_sphere), sizeof(rt_quad),
             sizeof(rt_material), sizeof(rt_texture), sizeof(rt_perlin), sizeof(rt_image), sizeof(rt_scene_desc),
             sizeof(rt_camera), sizeof(rt_opts), sizeof(rt_stats), sizeof(rt_scene_info));
      printf("%zu %zu %zu %zu\n", offsetof(rt_scene_desc, nodes), offsetof(rt_camera, pixel00),
             offsetof(rt_stats, kernel_ms), offsetof(rt_quad, normal));
      return 0;
    }"""
    with tempfile.TemporaryDirectory() as tmp:
        src = Path(tmp) / "sizes.c"
        src.write_text(prog)
        exe = Path(tmp) / "sizes"
        subprocess.run(["/usr/bin/gcc", "-std=c11", f"-I{ROOT / 'include'}", str(src), "-o", str(exe)], check=True)
        lines = subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split("\n")
    sizes = [int(v) for v in lines[0].split()]
    a = rtc.abi
    assert sizes == [a.NODE_DTYPE.itemsize, a.SPHERE_DTYPE.itemsize, a.QUAD_DTYPE.itemsize, a.MATERIAL_DTYPE.itemsize,
                     a.TEXTURE_DTYPE.itemsize, a.PERLIN_DTYPE.itemsize, a.IMAGE_DTYPE.itemsize,
                     C.sizeof(a.rt_scene_desc), C.sizeof(a.rt_camera), C.sizeof(a.rt_opts), C.sizeof(a.rt_stats),
                     C.sizeof(a.rt_scene_info)]
    offs = [int(v) for v in lines[1].split()]
    assert offs == [a.rt_scene_desc.nodes.offset, a.rt_camera.pixel00.offset, a.rt_stats.kernel_ms.offset,
                    a.QUAD_DTYPE.fields["normal"][1]]


@pytest.mark.parametrize("scene_id", range(8))
def test_inspect_builds_the_acceleration_structure_on_the_host(rtc, scene_factory, scene_id):
    _, desc = scene_factory(scene_id, 48, 4)
    info = rtc.inspect(desc)
    assert info["mega_supported"] == 1
    assert info["graph_frames"] <= 64 and info["graph_rays"] <= 6 and info["graph_recs"] <= 3
    assert info["bvh_depth"] + 2 <= 48
    n_media = int((desc.nodes["kind"] == rtc.abi.NODE_MEDIUM).sum())
    assert info["media"] == n_media
    if scene_id == 1:
        assert info["spheres"] == 486 and info["bvh_nodes"] < 486
    if scene_id == 6:
        assert info["quads"] == 18 and info["instances"] == 2 and info["chain_ops"] == 4
    if scene_id == 7:
        assert info["instances"] == 1 and info["media"] == 2 and info["quads"] == 2401


def _corrupt(desc, rtc, mutate):
    clone = rtc.abi.SceneDesc.from_bytes(desc.to_bytes())
    mutate(clone)
    return clone


def test_malformed_scenes_are_rejected(rtc, scene_factory):
    _, desc = scene_factory(6, 48, 4)
    info = rtc.abi.rt_scene_info()
    lib = rtc.lib.b200()

    def status(d):
        return lib.rt_scene_inspect(C.byref(d.c), C.byref(info))

    assert status(desc) == rtc.abi.RT_OK

    def bad_child(d):
        d.children[0] = 10_000

    def bad_material(d):
        d.quads["material"][0] = 99

    def bad_kind(d):
        d.nodes["kind"][3] = 42

    def cycle(d):
        lists = np.nonzero(d.nodes["kind"] == rtc.abi.NODE_LIST)[0]
        d.children[d.nodes["a"][lists[0]]] = lists[0]

    def bad_version(d):
        d.header.abi_version = 77

    for mutate in (bad_child, bad_material, bad_kind, cycle, bad_version):
        assert status(_corrupt(desc, rtc, mutate)) == rtc.abi.RT_ERR_INVALID, mutate.__name__
        assert lib.rt_last_error()


def test_lights_list_without_a_sampleable_item_is_rejected(rtc, scene_factory):
    _, desc = scene_factory(6, 48, 4)

    def strip(d):
        d.nodes["flags"][:] = np.where(d.nodes["kind"] == rtc.abi.NODE_LIST, d.nodes["flags"], 0)

    info = rtc.inspect(_corrupt(desc, rtc, strip))
    assert info["light_nesting"] == -1  # rt_scene_create turns this into RT_ERR_INVALID


def test_no_cpu_fallback(rtc, scene_factory):
    """Without a device every compute entry point must fail loudly, never compute on the host."""
    lib = rtc.lib.b200()
    if lib.rt_device_count() > 0:
        pytest.skip("a GPU is present")
    _, desc = scene_factory(0, 48, 4)
    with pytest.raises(rtc.lib.RtError) as err:
        rtc.Scene(desc)
    assert err.value.status == rtc.abi.RT_ERR_NO_DEVICE
    peak = C.c_float()
    assert lib.rt_fp32_peak(0, C.byref(peak)) == rtc.abi.RT_ERR_NO_DEVICE
    sess, _ = scene_factory(0, 48, 4)
    with pytest.raises(rtc.lib.RtError):
        sess.camera_render()


def test_scene_blob_round_trip(rtc, scene_factory):
    _, desc = scene_factory(7, 48, 4)
    again = rtc.abi.SceneDesc.from_bytes(desc.to_bytes())
    assert again.summary() == desc.summary()
    assert again.to_bytes() == desc.to_bytes()


def test_cli_matches_reference_argument_quirk(rtc):
    """`./main <scene> [width spp extra]`: width needs 3 arguments, spp needs 4 (src/main.c:289-292).  Without a GPU
    the CLI must abort (the reference's assert convention) rather than write an image from the host."""
    exe = rtc.lib.BUILD_DIR / "rtc_main"
    assert exe.exists()
    if rtc.lib.b200().rt_device_count() > 0:
        pytest.skip("a GPU is present; covered by the gpu tests")
    with tempfile.TemporaryDirectory() as tmp:
        proc = subprocess.run([str(exe), "0", "32", "1", "x"], cwd=tmp, capture_output=True, text=True)
        assert proc.returncode != 0
        assert "render failed" in proc.stderr
        assert not (Path(tmp) / "output.tiff").exists()


def test_flattened_instances_are_counted_and_capped(rtc, scene_factory):
    """Translate/RotateY chains do not become nodes of the rebuilt tree: their primitives join the group's one BVH
    tagged with the chain (8 bits).  60 chains: supported, 60 instances of 2 ops each; 300 chains: the throughput
    kernels refuse (RT_ERR_UNSUPPORTED at render time), the description stays valid for the exact kernel."""
    from synth_scenes import instanced_scene

    _, base = scene_factory(1, 48, 4)
    d60 = instanced_scene(rtc, base, 60)
    info = rtc.inspect(d60)
    assert info["mega_supported"] == 1
    assert info["instances"] == 60 and info["chain_ops"] == 120
    assert info["spheres"] + info["quads"] == 60 + 6 + 1
    assert info["bvh_nodes"] == info["spheres"] + info["quads"] - 1  # one primitive per leaf, one tree
    d300 = instanced_scene(rtc, base, 300)
    info = rtc.inspect(d300)
    assert info["mega_supported"] == 0 and info["instances"] >= 255


@pytest.mark.parametrize("name,shape", [("rgb_5x3", (5, 3, 3)), ("gray_4x2", (4, 2, 1)), ("rgb_300x2", (300, 2, 3))])
def test_write_tiff_bytes_equal_the_reference_writer(rtc, tmp_path, name, shape):
    """write_tiff (csrc/host/tiff_out.c) against files the reference's own writer produced
    (/root/reference/src/tiff.c:15-49 through `ref_tool tiff`, tests/golden/make_golden.py tiff): the same bytes,
    header, IFD, resolution rationals and strip, for RGB and for gray."""
    import ctypes as C

    from conftest import GOLDEN

    want = np.load(GOLDEN / "tiff.npz")[name].tobytes()
    w, h, c = shape
    n = w * h * c
    px = ((7 * np.arange(n, dtype=np.int64) + 3) % 251).astype(np.uint8)
    host = rtc.lib.host()
    libc = C.CDLL(None)
    libc.fopen.restype = C.c_void_p
    libc.fopen.argtypes = [C.c_char_p, C.c_char_p]
    libc.fclose.argtypes = [C.c_void_p]
    host.write_tiff.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
    path = tmp_path / "out.tiff"
    f = libc.fopen(str(path).encode(), b"wb")
    assert f
    assert host.write_tiff(f, w, h, c, px.ctypes.data_as(C.c_void_p)) == 0
    libc.fclose(f)
    assert path.read_bytes() == want


def test_builder_options_are_visible_to_inspect(rtc, scene_factory, monkeypatch):
    """rt_scene_inspect (host only) reports what the two optional builder steps did: oversized primitives taken out of
    the tree of a sphere-only world (the ground spheres of scenes 0 and 1; never in worlds with quads or media, never
    when fewer than two primitives would stay: the two equal spheres of scene 2, ground + marble ball of scene 4), and the 4-wide tree under
    RT_BVH_WIDE=1 (each wide node replaces one to three binary ones)."""
    want_pre = {0: 1, 1: 1, 2: 0, 4: 0, 5: 0, 6: 0, 7: 0}
    for scene_id, n_pre in want_pre.items():
        _, desc = scene_factory(scene_id, 48, 4)
        info = rtc.inspect(desc)
        assert info["pre_prims"] == n_pre, (scene_id, info)
        assert info["wide_nodes"] == 0
        plain_nodes = info["bvh_nodes"]
        monkeypatch.setenv("RT_BVH_PRE", "0")
        off = rtc.inspect(desc)
        monkeypatch.delenv("RT_BVH_PRE")
        assert off["pre_prims"] == 0 and off["bvh_nodes"] == plain_nodes + n_pre  # one leaf = one more binary node
        monkeypatch.setenv("RT_BVH_WIDE", "1")
        wide = rtc.inspect(desc)
        monkeypatch.delenv("RT_BVH_WIDE")
        if plain_nodes > 0:
            assert 0 < wide["wide_nodes"] <= plain_nodes and 3 * wide["wide_nodes"] >= plain_nodes
        assert wide["bvh_nodes"] == plain_nodes and wide["mega_supported"] == 1
