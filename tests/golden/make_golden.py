"""Generates tests/golden/scene<N>.npz from the UNMODIFIED reference (oracle/_ref/ref_tool, compiled from
/root/reference by oracle/Makefile with -O2 -ffp-contract=off).  Run in the build container, where /root/reference
exists:   python tests/golden/make_golden.py

Per scene the fixture holds what the reference itself produced:
  rgb8          Camera_render (src/raytracing.c:86) at width 48, 4 spp, depth 50 -- bytes of the output buffer
  camera_raw    the Camera struct after Camera_init (180 bytes)
  blob_sha256   sha256 of the flattened reference graph (this repo's flattener run over the reference's objects)
  rays_*        a fixed ray batch (camera rays, box-interior rays, axis-parallel edge cases) with per-ray PCG32 state
  hit_*         world.objects.hit() of the reference on that batch: t, p, normal, u, v, hit, front_face, material,
                advanced rng state
`python tests/golden/make_golden.py frames` writes tests/golden/frames.npz: the reference's own Camera_render
(src/raytracing.c:86-135, per-pixel seeds (17+j, 23+i)) at REAL geometry and sample counts -- Book 1 final 400x225 at
256 spp, Cornell box 200x200 at 256 spp, Book 2 final 200x200 at 128 spp, depth 50 -- as 8-bit frames, plus for the
noise floor the same frames from independent streams: the oracle in LIBM mode (pinned byte-identical to the reference
by test_oracle_golden.py) with the per-(pixel, sample) seeds.  The megakernel is tested against these on the GPU box.

The GPU box has no /root/reference; tests there pin the oracle (and the host layer) against these files.
"""
from __future__ import annotations

import hashlib
import importlib
import os
import subprocess
import sys
import tempfile
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
ROOT = HERE.parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
import raygen  # noqa: E402

rtc = importlib.import_module("ray-tracing-c_b200")
REF_TOOL = ROOT / "oracle" / "_ref" / "ref_tool"

RAY_IN = np.dtype([("o", "<f4", 3), ("d", "<f4", 3), ("t_min", "<f4"), ("t_max", "<f4"), ("state", "<u8"),
                   ("inc", "<u8")])
HIT_OUT = np.dtype([("t", "<f4"), ("p", "<f4", 3), ("normal", "<f4", 3), ("u", "<f4"), ("v", "<f4"), ("hit", "<u4"),
                    ("front", "<u4"), ("material", "<u4"), ("state", "<u8"), ("pad", "<u4", 2)])
assert RAY_IN.itemsize == 48 and HIT_OUT.itemsize == 64

WIDTH, SPP, DEPTH = 48, 4, 50
N_CAMERA, N_INTERIOR = 600, 600


def run(args, cwd):
    subprocess.run([str(REF_TOOL)] + [str(a) for a in args], cwd=cwd, check=True, stderr=subprocess.DEVNULL)


def ray_batch(scene_id):
    sess = rtc.Session(scene_id, width=WIDTH, spp=SPP)
    desc = sess.flatten()
    o1, d1 = raygen.camera_rays(sess.camera, N_CAMERA, seed=10 + scene_id)
    o2, d2 = raygen.secondary_rays(desc, N_INTERIOR, seed=20 + scene_id)
    oe, de = raygen.edge_case_rays()
    origins, dirs = np.concatenate([o1, o2, oe]), np.concatenate([d1, d2, de])
    rng = raygen.rng_states(len(origins), seed=30 + scene_id)
    return origins, dirs, rng


def main():
    if not REF_TOOL.exists():
        subprocess.run(["make", "-C", str(ROOT / "oracle"), "ref"], check=True)
    for scene_id in range(8):
        with tempfile.TemporaryDirectory() as tmp:
            run(["render", scene_id, WIDTH, SPP, DEPTH, 0, "out.rgb8", "camera.bin"], tmp)
            rgb8 = np.fromfile(os.path.join(tmp, "out.rgb8"), dtype=np.uint8)
            camera_raw = np.fromfile(os.path.join(tmp, "camera.bin"), dtype=np.uint8)
            run(["flatten", scene_id, "scene.blob"], tmp)
            blob = Path(tmp, "scene.blob").read_bytes()

            origins, dirs, rng = ray_batch(scene_id)
            rays = np.zeros(len(origins), dtype=RAY_IN)
            rays["o"], rays["d"] = origins, dirs
            rays["t_min"], rays["t_max"] = 1e-3, np.inf
            rays["state"], rays["inc"] = rng[:, 0], rng[:, 1]
            rays.tofile(os.path.join(tmp, "rays.bin"))
            run(["hits", scene_id, "rays.bin", "hits.bin"], tmp)
            hits = np.fromfile(os.path.join(tmp, "hits.bin"), dtype=HIT_OUT)

        height = len(rgb8) // (WIDTH * 3)
        out = HERE / f"scene{scene_id}.npz"
        np.savez_compressed(
            out, rgb8=rgb8.reshape(height, WIDTH, 3), camera_raw=camera_raw,
            blob_sha256=np.frombuffer(hashlib.sha256(blob).digest(), dtype=np.uint8),
            width=WIDTH, spp=SPP, depth=DEPTH,
            rays_o=origins, rays_d=dirs, rays_rng=rng,
            hit_t=hits["t"], hit_p=hits["p"], hit_normal=hits["normal"], hit_u=hits["u"], hit_v=hits["v"],
            hit_hit=hits["hit"], hit_front=hits["front"], hit_material=hits["material"], hit_state=hits["state"])
        print(f"scene {scene_id}: {WIDTH}x{height}, {len(origins)} rays, {int(hits['hit'].sum())} hits, "
              f"blob {len(blob)} B -> {out.name} ({out.stat().st_size} B)")


FRAMES = [  # scene, width, spp: BASELINE.json configs 2, 3, 4 at a third / a fifth of their linear size
    (1, 400, 256),
    (6, 200, 256),
    (7, 200, 128),
]


def frames():
    import oracle_lib as O

    if not REF_TOOL.exists():
        subprocess.run(["make", "-C", str(ROOT / "oracle"), "ref"], check=True)
    out = {}
    for scene_id, width, spp in FRAMES:
        with tempfile.TemporaryDirectory() as tmp:
            run(["render", scene_id, width, spp, DEPTH, 0, "out.rgb8", "camera.bin"], tmp)
            rgb8 = np.fromfile(os.path.join(tmp, "out.rgb8"), dtype=np.uint8)
        height = len(rgb8) // (width * 3)
        ref = rgb8.reshape(height, width, 3)
        sess = rtc.Session(scene_id, width=width, spp=spp, depth=DEPTH)
        assert (sess.height, sess.width) == (height, width)
        other, _, _ = O.render(sess.flatten(), sess.camera, rtc.make_opts(kernel="exact", rng="sample"), O.MATH_LIBM)
        d = ref.astype(np.float64) - other.astype(np.float64)
        print(f"scene {scene_id}: {width}x{height}x{spp} spp; reference vs independent streams: RMSE "
              f"{np.sqrt((d ** 2).mean(axis=(0, 1))) } /255 per channel, mean diff {d.mean(axis=(0, 1))}")
        out[f"ref_{scene_id}"] = ref
        out[f"other_{scene_id}"] = other
        out[f"shape_{scene_id}"] = np.array([width, height, spp, DEPTH])
    path = HERE / "frames.npz"
    np.savez_compressed(path, **out)
    print(f"-> {path.name} ({path.stat().st_size} B)")


def tiffs():
    """tests/golden/tiff.npz: files written by the reference's own write_tiff (src/tiff.c:15-49), RGB and gray."""
    out = {}
    for name, (w, h, c) in {"rgb_5x3": (5, 3, 3), "gray_4x2": (4, 2, 1), "rgb_300x2": (300, 2, 3)}.items():
        with tempfile.TemporaryDirectory() as tmp:
            run(["tiff", 0, w, h, c, "out.tiff"], tmp)
            out[name] = np.frombuffer(Path(tmp, "out.tiff").read_bytes(), dtype=np.uint8)
    path = HERE / "tiff.npz"
    np.savez_compressed(path, **out)
    print(f"-> {path.name} ({path.stat().st_size} B)")


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "scenes"
    if what in ("tiff", "all"):
        tiffs()
    if what in ("scenes", "all"):
        main()
    if what in ("frames", "all"):
        frames()
