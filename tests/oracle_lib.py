"""ctypes access to the CPU oracle (oracle/_build/liboracle.so) -- TEST INFRASTRUCTURE, never imported by the
product package."""
from __future__ import annotations

import ctypes as C
import importlib
import subprocess
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
ORACLE_DIR = ROOT / "oracle"
MATH_LIBM, MATH_PORTABLE = 0, 1

rtc = importlib.import_module("ray-tracing-c_b200")
abi = rtc.abi

_lib = None


def build_oracle():
    subprocess.run(["make", "-C", str(ORACLE_DIR), "all"], check=True, capture_output=True)


def oracle() -> C.CDLL:
    global _lib
    if _lib is None:
        path = ORACLE_DIR / "_build" / "liboracle.so"
        if not path.exists():
            build_oracle()
        lib = C.CDLL(str(path))
        vp = C.c_void_p
        lib.oracle_render.argtypes = [C.POINTER(abi.rt_scene_desc), C.POINTER(abi.rt_camera), C.POINTER(abi.rt_opts),
                                      C.c_int, vp, vp, C.POINTER(abi.rt_stats)]
        lib.oracle_hit_batch.argtypes = [C.POINTER(abi.rt_scene_desc), C.c_int, C.c_size_t, vp, vp, C.c_float,
                                         C.c_float, vp, vp, vp, vp, vp, vp, vp, vp]
        lib.oracle_tonemap.argtypes = [vp, C.c_size_t, C.c_int, vp]
        lib.oracle_tonemap.restype = None
        lib.oracle_pcg32.argtypes = [C.c_uint64, C.c_uint64, C.c_size_t, vp]
        lib.oracle_pcg32.restype = None
        lib.oracle_pm.argtypes = [C.c_int, C.c_float, C.c_float]
        lib.oracle_pm.restype = C.c_float
        _lib = lib
    return _lib


def render(desc, camera, opts, math=MATH_PORTABLE):
    h, w = camera.img_height, camera.img_width
    rad = np.empty((h, w, 3), np.float32)
    rgb8 = np.empty((h, w, 3), np.uint8)
    stats = abi.rt_stats()
    rc = oracle().oracle_render(C.byref(desc.c), C.byref(camera), C.byref(opts), math, rad.ctypes.data,
                                rgb8.ctypes.data, C.byref(stats))
    assert rc == 0, rc
    return rgb8, rad, stats.as_dict()


def hit_batch(desc, origins, dirs, t_min=1e-3, t_max=np.inf, rng=None, math=MATH_PORTABLE):
    origins = np.ascontiguousarray(origins, np.float32).reshape(-1, 3)
    dirs = np.ascontiguousarray(dirs, np.float32).reshape(-1, 3)
    n = origins.shape[0]
    out = {
        "t": np.empty(n, np.float32), "node": np.empty(n, np.uint32), "p": np.zeros((n, 3), np.float32),
        "normal": np.zeros((n, 3), np.float32), "uv": np.zeros((n, 2), np.float32),
        "front": np.zeros(n, np.uint8), "material": np.empty(n, np.uint32),
    }
    rng_io = None
    if rng is not None:
        rng_io = np.ascontiguousarray(rng, np.uint64).reshape(n, 2).copy()
        out["rng"] = rng_io
    rc = oracle().oracle_hit_batch(C.byref(desc.c), math, n, origins.ctypes.data, dirs.ctypes.data, float(t_min),
                                   float(t_max), rng_io.ctypes.data if rng_io is not None else None,
                                   out["t"].ctypes.data, out["node"].ctypes.data, out["p"].ctypes.data,
                                   out["normal"].ctypes.data, out["uv"].ctypes.data, out["front"].ctypes.data,
                                   out["material"].ctypes.data)
    assert rc == 0, rc
    return out


def tonemap(radiance_sum, spp):
    flat = np.ascontiguousarray(radiance_sum, np.float32).reshape(-1)
    out = np.empty(flat.shape, np.uint8)
    oracle().oracle_tonemap(flat.ctypes.data, flat.size, spp, out.ctypes.data)
    return out.reshape(np.shape(radiance_sum))


def pcg32(initstate, initseq, n):
    out = np.empty(n, np.uint32)
    oracle().oracle_pcg32(initstate, initseq, n, out.ctypes.data)
    return out


# the algorithmic-work model of SURVEY.md section 8(d): FLOP per visit of the REFERENCE's own traversal
FLOP_AABB, FLOP_SPHERE, FLOP_QUAD, FLOP_MEDIUM, FLOP_XFORM, FLOP_FINALISE, FLOP_SCATTER = 27, 24, 36, 30, 14, 40, 60


def last_work() -> dict:
    """Visits of the last render(): AABB / sphere / quad tests, medium evaluations, instance transforms, segments."""
    out = (C.c_uint64 * 6)()
    oracle().oracle_last_work(out)
    keys = ("aabb", "sphere", "quad", "medium", "xform", "segments")
    return dict(zip(keys, [int(v) for v in out]))


def flop_per_segment(work: dict) -> float:
    seg = max(work["segments"], 1)
    return (FLOP_AABB * work["aabb"] + FLOP_SPHERE * work["sphere"] + FLOP_QUAD * work["quad"] +
            FLOP_MEDIUM * work["medium"] + FLOP_XFORM * work["xform"]) / seg + FLOP_FINALISE + FLOP_SCATTER
