"""ctypes bindings of the two in-tree shared libraries.

  librtc_b200.so  the sm_100a CUDA kernels behind the C ABI of include/rtc_b200.h   (csrc/device)
  librtc_host.so  the reference's C API surface + flattener + scene builders        (csrc/host)

There is no fallback of any kind: a missing library raises ImportError-like RuntimeError with the build command,
and every compute entry point returns RT_ERR_NO_DEVICE on a machine without a GPU.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

from . import abi

PKG_DIR = Path(__file__).resolve().parent
# RTC_BUILD_DIR: a second build of the same sources (make OUT=... EXTRA=-D...) for A/B measurements
BUILD_DIR = Path(os.environ.get("RTC_BUILD_DIR") or PKG_DIR / "_build")
REPO_ROOT = PKG_DIR.parent

# every symbol include/rtc_b200.h declares (tests check the .so exports each one)
B200_SYMBOLS = [
    "rt_last_error", "rt_abi_version", "rt_device_count", "rt_scene_inspect", "rt_scene_create", "rt_scene_destroy",
    "rt_scene_acquire", "rt_scene_release", "rt_shutdown", "rt_render",
    "rt_render_device", "rt_render_device_noalloc", "rt_render_device_stats", "rt_tonemap_device", "rt_tonemap", "rt_hit_batch", "rt_render_multi",
    "rt_fp32_peak",
]
HOST_SYMBOLS = [
    # the reference's API (include/rtc_api.h)
    "Camera_init", "Camera_render", "World_init", "write_tiff", "my_malloc", "pcg32_seed", "pcg32_u32", "pcg32_f32", "pcg32_advance",
    "HittableList_new", "HittableList_append", "Sphere_new", "Quad_new", "Box_new", "BVHNode_new", "Translate_new",
    "RotateY_new", "ConstantMedium_new", "Lambertian_new", "Metal_new", "Dielectric_new", "DiffuseLight_new",
    "Isotropic_new", "Solid_new", "Checker_new", "Image_new", "Perlin_new",
    # helpers (include/rtc_host.h)
    "rtc_flatten", "rtc_flatten_error", "rtc_scene_desc_free", "rtc_scene_desc_write", "rtc_scene_desc_read",
    "rtc_camera_pack", "rtc_camera_defaults", "rtc_scene_build", "rtc_scene_title", "rtc_camera_render_opts",
    "rtc_session_open", "rtc_session_close", "rtc_session_width", "rtc_session_height", "rtc_session_spp",
    "rtc_session_camera", "rtc_session_camera_raw", "rtc_session_flatten", "rtc_session_write_blob",
    "rtc_session_render", "rtc_session_camera_render",
]


class BackendMissing(RuntimeError):
    pass


def build(verbose: bool = False) -> None:
    """Compile both libraries in-tree for sm_100a (nvcc cross-compiles without a GPU)."""
    import subprocess

    cmd = ["make", "-C", str(PKG_DIR / "csrc"), "-j8", "all"]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or proc.returncode != 0:
        print(proc.stdout[-4000:])
        print(proc.stderr[-4000:])
    if proc.returncode != 0:
        raise RuntimeError("building librtc_b200.so / librtc_host.so failed")


_b200 = None
_host = None


def _load(name: str) -> C.CDLL:
    path = BUILD_DIR / name
    if not path.exists():
        raise BackendMissing(
            f"{path} is missing: build it with `make -C {PKG_DIR / 'csrc'}` (or __graft_entry__.build()). "
            "There is no CPU fallback."
        )
    return C.CDLL(str(path), mode=C.RTLD_GLOBAL)


def _point_at_bundled_nccl() -> None:
    """rt_render_multi dlopens NCCL at its first multi-GPU call.  In a Python process that may import torch LATER the
    library under the soname libnccl.so.2 must be the one torch was built against (the wheel bundles it under
    nvidia/nccl/lib), not an older system copy: tell the C side where it is (RT_NCCL_LIB), without importing torch."""
    if os.environ.get("RT_NCCL_LIB"):
        return
    import importlib.util

    try:
        spec = importlib.util.find_spec("nvidia")
    except (ImportError, ValueError):
        spec = None
    for base in (spec.submodule_search_locations if spec and spec.submodule_search_locations else []):
        cand = Path(base) / "nccl" / "lib" / "libnccl.so.2"
        if cand.exists():
            os.environ["RT_NCCL_LIB"] = str(cand)
            return


def b200() -> C.CDLL:
    """librtc_b200.so with argtypes declared."""
    global _b200
    if _b200 is not None:
        return _b200
    _point_at_bundled_nccl()
    lib = _load("librtc_b200.so")
    vp, u32, i32, sz, f32 = C.c_void_p, C.c_uint32, C.c_int, C.c_size_t, C.c_float
    lib.rt_last_error.restype = C.c_char_p
    lib.rt_abi_version.restype = u32
    lib.rt_device_count.restype = i32
    lib.rt_scene_create.argtypes = [C.POINTER(abi.rt_scene_desc), i32, C.POINTER(vp)]
    lib.rt_scene_destroy.argtypes = [vp]
    lib.rt_scene_acquire.argtypes = [C.POINTER(abi.rt_scene_desc), i32, C.POINTER(vp)]
    lib.rt_scene_release.argtypes = [vp]
    lib.rt_scene_release.restype = None
    lib.rt_shutdown.restype = None
    lib.rt_scene_inspect.argtypes = [C.POINTER(abi.rt_scene_desc), C.POINTER(abi.rt_scene_info)]
    lib.rt_scene_destroy.restype = None
    lib.rt_render.argtypes = [vp, C.POINTER(abi.rt_camera), C.POINTER(abi.rt_opts), vp, vp, C.POINTER(abi.rt_stats)]
    lib.rt_render_device.argtypes = [vp, C.POINTER(abi.rt_camera), C.POINTER(abi.rt_opts), vp, vp]
    lib.rt_render_device_noalloc.argtypes = [vp, C.POINTER(abi.rt_camera), C.POINTER(abi.rt_opts), vp, vp]
    lib.rt_render_device_stats.argtypes = [vp, vp, C.POINTER(abi.rt_stats)]
    lib.rt_tonemap_device.argtypes = [vp, sz, i32, vp, vp]
    lib.rt_tonemap.argtypes = [vp, sz, i32, vp]
    lib.rt_hit_batch.argtypes = [vp, u32, sz, vp, vp, f32, f32, vp, vp, vp, vp, vp, vp, vp, vp]
    lib.rt_render_multi.argtypes = [C.POINTER(abi.rt_scene_desc), C.POINTER(abi.rt_camera), C.POINTER(abi.rt_opts),
                                    i32, vp, vp, C.POINTER(abi.rt_stats)]
    lib.rt_fp32_peak.argtypes = [i32, C.POINTER(f32)]
    _b200 = lib
    return lib


def host() -> C.CDLL:
    """librtc_host.so (pulls librtc_b200.so in through its rpath)."""
    global _host
    if _host is not None:
        return _host
    b200()
    lib = _load("librtc_host.so")
    vp, i32, f32 = C.c_void_p, C.c_int, C.c_float
    lib.rtc_session_open.argtypes = [i32, i32, i32, i32, f32, f32]
    lib.rtc_session_open.restype = vp
    lib.rtc_session_close.argtypes = [vp]
    lib.rtc_session_close.restype = None
    for name in ("rtc_session_width", "rtc_session_height", "rtc_session_spp"):
        getattr(lib, name).argtypes = [vp]
    lib.rtc_session_camera.argtypes = [vp, C.POINTER(abi.rt_camera)]
    lib.rtc_session_camera.restype = None
    lib.rtc_session_camera_raw.argtypes = [vp, vp, C.c_size_t]
    lib.rtc_session_camera_raw.restype = C.c_size_t
    lib.rtc_session_flatten.argtypes = [vp, C.POINTER(abi.rt_scene_desc)]
    lib.rtc_session_write_blob.argtypes = [vp, C.c_char_p]
    lib.rtc_session_render.argtypes = [vp, C.POINTER(abi.rt_opts), i32, vp, vp, C.POINTER(abi.rt_stats)]
    lib.rtc_session_camera_render.argtypes = [vp, vp]
    lib.rtc_session_camera_render.restype = None
    lib.rtc_scene_desc_free.argtypes = [C.POINTER(abi.rt_scene_desc)]
    lib.rtc_scene_desc_free.restype = None
    lib.rtc_flatten_error.restype = C.c_char_p
    lib.rtc_synthetic_image.argtypes = [i32, i32, vp]
    lib.rtc_synthetic_image.restype = None
    lib.rtc_scene_title.argtypes = [i32]
    lib.rtc_scene_title.restype = C.c_char_p
    _host = lib
    return lib


class RtError(RuntimeError):
    def __init__(self, status: int, where: str):
        msg = b200().rt_last_error()
        super().__init__(f"{where} failed with rt_status {status}: {msg.decode() if msg else ''}")
        self.status = status


def check(status: int, where: str) -> None:
    if status != abi.RT_OK:
        raise RtError(status, where)
