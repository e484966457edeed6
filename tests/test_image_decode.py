"""Image-texture ingest: csrc/host/image_decode.c against stb_image, the decoder behind the
reference's Image_init (/root/reference/src/texture.c:38-42, stbi_load(..., 3)).  tests/golden/images/ holds small PNG
and JPEG files and stb.npz what the REFERENCE decodes from them (tests/golden/make_golden.py images, through
`oracle/_ref/ref_tool image`).  Bar: the same bytes -- for JPEG that pins the inverse DCT, the chroma upsampling and the
colour conversion, not just "a plausible picture"."""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import GOLDEN, ROOT

IMAGES = GOLDEN / "images"
NAMES = sorted(p.name for p in IMAGES.iterdir() if p.suffix in (".png", ".jpg"))


def _load(rtc, path):
    host = rtc.lib.host()
    host.rtc_load_image.restype = C.POINTER(C.c_uint8)
    host.rtc_load_image.argtypes = [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]
    host.rtc_image_error.restype = C.c_char_p
    w, h, exists = C.c_int(), C.c_int(), C.c_int()
    p = host.rtc_load_image(str(path).encode(), C.byref(w), C.byref(h), C.byref(exists))
    if not p:
        return None, bool(exists.value), host.rtc_image_error().decode()
    return np.ctypeslib.as_array(p, shape=(h.value, w.value, 3)).copy(), True, ""


@pytest.mark.parametrize("name", [n for n in NAMES if "progressive" not in n])
def test_decoder_gives_the_bytes_stb_gives(rtc, name):
    want = np.load(IMAGES / "stb.npz")[name]
    got, exists, why = _load(rtc, IMAGES / name)
    assert got is not None, why
    assert got.shape == want.shape
    assert np.array_equal(got, want)


def test_what_is_not_decoded_fails_loudly(rtc, tmp_path):
    got, exists, why = _load(rtc, IMAGES / "a_progressive.jpg")
    assert got is None and exists and "progressive" in why
    got, exists, why = _load(rtc, tmp_path / "no_such_file.jpg")
    assert got is None and not exists
    junk = tmp_path / "junk.jpg"
    junk.write_bytes(b"\xff\xd8\xff\xdb\x00\x05garbage")
    got, exists, why = _load(rtc, junk)
    assert got is None and exists and why
    truncated = tmp_path / "cut.png"
    truncated.write_bytes((IMAGES / "b_rgb.png").read_bytes()[:200])
    got, exists, why = _load(rtc, truncated)
    assert got is None and exists and why


def test_image_texture_uses_the_file_in_the_working_directory(rtc, tmp_path):
    """Scene 3 (`Image_new("earthmap.jpg")`, src/main.c:104): with a real JPEG in the working directory the flattened
    scene carries its texels (as stb decodes them); with an undecodable file of that name the program stops like the
    reference's assert (texture.c:41); only without the file does the synthetic image stand in."""
    code = ("import sys, importlib, numpy as np; sys.path.insert(0, %r); rtc = importlib.import_module('ray-tracing-c_b200');"
            "s = rtc.Session(3, width=32, spp=1); d = s.flatten(); print(int(d.images['width'][0]), int(d.images['height'][0]),"
            " int(np.asarray(d.texels, dtype=np.uint64).sum()))" % str(ROOT))
    env = dict(os.environ)
    (tmp_path / "earthmap.jpg").write_bytes((IMAGES / "a_420_q50.jpg").read_bytes())
    out = subprocess.run([sys.executable, "-c", code], cwd=tmp_path, capture_output=True, text=True, env=env)
    assert out.returncode == 0, out.stderr
    want = np.load(IMAGES / "stb.npz")["a_420_q50.jpg"]
    w, h, total = (int(v) for v in out.stdout.split())
    assert (w, h) == (want.shape[1], want.shape[0]) and total == int(want.astype(np.uint64).sum())
    assert "synthetic" not in out.stderr
    (tmp_path / "earthmap.jpg").write_bytes((IMAGES / "a_progressive.jpg").read_bytes())
    out = subprocess.run([sys.executable, "-c", code], cwd=tmp_path, capture_output=True, text=True, env=env)
    assert out.returncode != 0 and "cannot decode image" in out.stderr
    (tmp_path / "earthmap.jpg").unlink()
    out = subprocess.run([sys.executable, "-c", code], cwd=tmp_path, capture_output=True, text=True, env=env)
    assert out.returncode == 0 and "synthetic" in out.stderr
    assert out.stdout.split()[:2] == ["1024", "512"]
