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


def _natural(w, h, seed):
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:h, 0:w]
    img = np.stack([128 + 100 * np.sin(x / 17.0 + y / 29.0), 128 + 90 * np.cos(x / 11.0 - y / 7.0), (x * 3 + y * 5) % 256], -1)
    img = img + rng.normal(0, 12, img.shape)
    img[h // 3:h // 3 + 5, w // 4:w // 4 + 9] = [255, 0, 40]
    return np.clip(img, 0, 255).astype(np.uint8)


def _adam7_png(rgb):
    """An interlaced (Adam7) 8-bit RGB PNG, filter type 0 on every row (PIL cannot write interlaced files)."""
    import struct
    import zlib

    h, w, _ = rgb.shape
    raw = b""
    for x0, y0, dx, dy in ((0, 0, 8, 8), (4, 0, 8, 8), (0, 4, 4, 8), (2, 0, 4, 4), (0, 2, 2, 4), (1, 0, 2, 2), (0, 1, 1, 2)):
        sub = rgb[y0::dy, x0::dx]
        if sub.shape[0] == 0 or sub.shape[1] == 0:
            continue
        for row in sub:
            raw += b"\x00" + row.tobytes()

    def chunk(kind, body):
        return struct.pack(">I", len(body)) + kind + body + struct.pack(">I", zlib.crc32(kind + body))

    return (b"\x89PNG\r\n\x1a\n" + chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, 8, 2, 0, 0, 1)) +
            chunk(b"IDAT", zlib.compress(raw, 6)) + chunk(b"IEND", b""))


def images():
    """tests/golden/images/: small PNG and JPEG files (made with PIL) and, in stb.npz, what the reference's own
    Image_init (src/texture.c:38-42: stbi_load(..., 3)) decodes from each -- through `ref_tool image`."""
    from PIL import Image

    out_dir = HERE / "images"
    out_dir.mkdir(exist_ok=True)
    for old in out_dir.iterdir():
        old.unlink()
    a, b = Image.fromarray(_natural(61, 37, 1)), Image.fromarray(_natural(48, 32, 2))
    a.save(out_dir / "a_444_q90.jpg", quality=90, subsampling=0)
    a.save(out_dir / "a_422_q75.jpg", quality=75, subsampling=1)
    a.save(out_dir / "a_420_q50.jpg", quality=50, subsampling=2)
    a.save(out_dir / "a_420_rst.jpg", quality=85, subsampling=2, restart_marker_blocks=2)
    a.save(out_dir / "a_opt.jpg", quality=85, optimize=True)
    a.convert("L").save(out_dir / "a_gray.jpg", quality=80)
    Image.fromarray(_natural(1, 1, 3)).save(out_dir / "one_pixel.jpg", quality=90)
    a.save(out_dir / "a_progressive.jpg", quality=80, progressive=True)  # decoded by stb, refused loudly here
    b.save(out_dir / "b_rgb.png")
    b.convert("RGBA").save(out_dir / "b_rgba.png")
    b.convert("L").save(out_dir / "b_gray.png")
    b.convert("LA").save(out_dir / "b_gray_alpha.png")
    b.convert("P", palette=Image.ADAPTIVE, colors=17).save(out_dir / "b_palette.png")
    b.convert("1").save(out_dir / "b_1bit.png")
    Image.fromarray(_natural(48, 32, 2)[..., 0].astype(np.uint16) * 257 + 13).save(out_dir / "b_gray16.png")
    (out_dir / "b_adam7.png").write_bytes(_adam7_png(_natural(21, 13, 4)))
    decoded = {}
    for path in sorted(out_dir.iterdir()):
        with tempfile.TemporaryDirectory() as tmp:
            run(["image", 0, path, "out.bin"], tmp)
            raw = Path(tmp, "out.bin").read_bytes()
        nl = raw.index(b"\n")
        w, h = (int(v) for v in raw[:nl].split())
        decoded[path.name] = np.frombuffer(raw[nl + 1:], np.uint8).reshape(h, w, 3)
    np.savez_compressed(HERE / "images" / "stb.npz", **decoded)
    print(f"-> images/ ({len(decoded)} files + stb.npz)")


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "scenes"
    if what in ("tiff", "all"):
        tiffs()
    if what in ("images", "all"):
        images()
    if what in ("scenes", "all"):
        main()
    if what in ("frames", "all"):
        frames()
