"""Host-side mirror of the reference's render interface, in Python, over the C ABI.

  Session   one built-in scene (`./main <scene> [width spp]` of /root/reference/src/main.c:275-345): World + Camera
            built by the C host library; `.camera_render()` is Camera_render (raytracing.c:86) end to end with HOST
            buffers; `.flatten()` gives the rt_scene_desc the kernels consume.
  Scene     a flattened scene resident on one GPU: `.render()` (rt_render), `.render_device()` (rt_render_device into
            a torch tensor, for the one-process-per-GPU driver in dist.py) and `.hit_batch()` (rt_hit_batch).

numpy holds host buffers; torch is only used (optionally) for device memory and streams.
"""
from __future__ import annotations

import ctypes as C
import tempfile
from pathlib import Path

import numpy as np

from . import abi, lib


def make_opts(kernel="mega", rng="sample", sample_begin=0, sample_count=0, drop_nan=False) -> abi.rt_opts:
    o = abi.rt_opts()
    o.kernel = abi.KERNELS[kernel] if isinstance(kernel, str) else int(kernel)
    o.rng_mode = abi.RNG_MODES[rng] if isinstance(rng, str) else int(rng)
    o.sample_begin = int(sample_begin)
    o.sample_count = int(sample_count)
    o.flags = abi.FLAG_DROP_NAN if drop_nan else 0
    return o


class Session:
    """A scene of the reference's CLI, built on the host by the drop-in constructors."""

    def __init__(self, scene_id: int, width: int = 0, spp: int = 0, depth: int = 0, aspect: float = -1.0,
                 lights_prob: float = -1.0):
        self._h = lib.host()
        self._p = self._h.rtc_session_open(scene_id, width, spp, depth, aspect, lights_prob)
        if not self._p:
            raise ValueError(f"unknown scene {scene_id}")
        self.scene_id = scene_id
        self.width = self._h.rtc_session_width(self._p)
        self.height = self._h.rtc_session_height(self._p)
        self.spp = self._h.rtc_session_spp(self._p)

    def close(self):
        if self._p:
            self._h.rtc_session_close(self._p)
            self._p = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def camera(self) -> abi.rt_camera:
        cam = abi.rt_camera()
        self._h.rtc_session_camera(self._p, C.byref(cam))
        return cam

    def camera_raw(self) -> bytes:
        buf = C.create_string_buffer(256)
        n = self._h.rtc_session_camera_raw(self._p, buf, 256)
        return buf.raw[:n]

    def flatten(self) -> abi.SceneDesc:
        """rtc_flatten through the blob format (so the arrays live in numpy, independent of the C heap)."""
        with tempfile.TemporaryDirectory() as tmp:
            path = str(Path(tmp) / "scene.blob")
            if self._h.rtc_session_write_blob(self._p, path.encode()) != 0:
                raise RuntimeError("flatten failed: " + self._h.rtc_flatten_error().decode())
            return abi.SceneDesc.load(path)

    def camera_render(self, opts: abi.rt_opts | None = None, n_gpus: int = 1, want_radiance: bool = False):
        """Camera_render with explicit options: World (host pointers) -> rgb8 (host).  Returns (rgb8, radiance|None,
        stats dict).  Includes flatten + BVH build + upload + kernels + download, like the reference's timed call."""
        opts = opts or make_opts()
        rgb8 = np.empty((self.height, self.width, 3), dtype=np.uint8)
        rad = np.empty((self.height, self.width, 3), dtype=np.float32) if want_radiance else None
        stats = abi.rt_stats()
        rc = self._h.rtc_session_render(self._p, C.byref(opts), n_gpus, rgb8.ctypes.data,
                                        rad.ctypes.data if rad is not None else None, C.byref(stats))
        lib.check(rc, "rtc_session_render")
        return rgb8, rad, stats.as_dict()


class Scene:
    """A flattened scene uploaded to one GPU (rt_scene_create)."""

    def __init__(self, desc: abi.SceneDesc, device: int = -1, cached: bool = False):
        """cached=True: rt_scene_acquire / rt_scene_release -- the uploaded scene comes from (and goes back to) the
        library's cache keyed by the description's content, as Camera_render does; a second Scene of an equal
        description skips the BVH build and the upload."""
        self._l = lib.b200()
        self.desc = desc
        self._p = C.c_void_p()
        self._cached = cached
        if cached:
            lib.check(self._l.rt_scene_acquire(C.byref(desc.c), device, C.byref(self._p)), "rt_scene_acquire")
        else:
            lib.check(self._l.rt_scene_create(C.byref(desc.c), device, C.byref(self._p)), "rt_scene_create")

    def close(self):
        if self._p:
            if self._cached:
                self._l.rt_scene_release(self._p)
            else:
                self._l.rt_scene_destroy(self._p)
            self._p = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def render(self, camera: abi.rt_camera, opts: abi.rt_opts | None = None, want_radiance=True, want_rgb8=True):
        opts = opts or make_opts()
        h, w = camera.img_height, camera.img_width
        rgb8 = np.empty((h, w, 3), dtype=np.uint8) if want_rgb8 else None
        rad = np.empty((h, w, 3), dtype=np.float32) if want_radiance else None
        stats = abi.rt_stats()
        rc = self._l.rt_render(self._p, C.byref(camera), C.byref(opts), rad.ctypes.data if rad is not None else None,
                               rgb8.ctypes.data if rgb8 is not None else None, C.byref(stats))
        lib.check(rc, "rt_render")
        return rgb8, rad, stats.as_dict()

    def render_device(self, camera: abi.rt_camera, opts: abi.rt_opts, radiance_ptr: int, stream_ptr: int = 0):
        """Asynchronous render into device memory (float[h][w][3]) on a CUDA stream."""
        rc = self._l.rt_render_device(self._p, C.byref(camera), C.byref(opts), radiance_ptr, stream_ptr or None)
        lib.check(rc, "rt_render_device")

    def render_device_noalloc(self, camera: abi.rt_camera, opts: abi.rt_opts, radiance_ptr: int, stream_ptr: int = 0):
        """rt_render_device_noalloc: the same render with its scratch in the library's static device array -- no
        allocation in the call; frames up to RT_NOALLOC_MAX_PIXELS; calls that overlap on the GPU race."""
        rc = self._l.rt_render_device_noalloc(self._p, C.byref(camera), C.byref(opts), radiance_ptr, stream_ptr or None)
        lib.check(rc, "rt_render_device_noalloc")

    def device_stats(self, stream_ptr: int = 0) -> dict:
        stats = abi.rt_stats()
        lib.check(self._l.rt_render_device_stats(self._p, stream_ptr or None, C.byref(stats)), "rt_render_device_stats")
        return stats.as_dict()

    def hit_batch(self, origins, dirs, t_min=1e-3, t_max=np.inf, rng=None, kernel="exact"):
        """Closest hits of a ray batch (the HittableVTable.hit slot, include/hittable.h:30).  Returns a dict of numpy
        arrays: t, node, p, normal, uv, front, material (+ rng with the advanced states if `rng` was given)."""
        origins = np.ascontiguousarray(origins, dtype=np.float32).reshape(-1, 3)
        dirs = np.ascontiguousarray(dirs, dtype=np.float32).reshape(-1, 3)
        n = origins.shape[0]
        assert dirs.shape[0] == n
        out = {
            "t": np.empty(n, np.float32), "node": np.empty(n, np.uint32), "p": np.zeros((n, 3), np.float32),
            "normal": np.zeros((n, 3), np.float32), "uv": np.zeros((n, 2), np.float32),
            "front": np.zeros(n, np.uint8), "material": np.empty(n, np.uint32),
        }
        rng_io = None
        if rng is not None:
            rng_io = np.ascontiguousarray(rng, dtype=np.uint64).reshape(n, 2).copy()
            out["rng"] = rng_io
        k = abi.KERNELS[kernel] if isinstance(kernel, str) else int(kernel)
        rc = self._l.rt_hit_batch(self._p, k, n, origins.ctypes.data, dirs.ctypes.data, float(t_min), float(t_max),
                                  rng_io.ctypes.data if rng_io is not None else None, out["t"].ctypes.data,
                                  out["node"].ctypes.data, out["p"].ctypes.data, out["normal"].ctypes.data,
                                  out["uv"].ctypes.data, out["front"].ctypes.data, out["material"].ctypes.data)
        lib.check(rc, "rt_hit_batch")
        return out


def render_multi(desc: abi.SceneDesc, camera: abi.rt_camera, opts: abi.rt_opts, n_gpus: int, want_radiance=True):
    """rt_render_multi: one process drives n GPUs, sample-range shards, peer-copy gather on device 0."""
    l = lib.b200()
    h, w = camera.img_height, camera.img_width
    rgb8 = np.empty((h, w, 3), dtype=np.uint8)
    rad = np.empty((h, w, 3), dtype=np.float32) if want_radiance else None
    stats = abi.rt_stats()
    rc = l.rt_render_multi(C.byref(desc.c), C.byref(camera), C.byref(opts), n_gpus,
                           rad.ctypes.data if rad is not None else None, rgb8.ctypes.data, C.byref(stats))
    lib.check(rc, "rt_render_multi")
    return rgb8, rad, stats.as_dict()


def inspect(desc: abi.SceneDesc) -> dict:
    """rt_scene_inspect: host-only validation + shape of the rebuilt acceleration structure (no GPU needed)."""
    info = abi.rt_scene_info()
    lib.check(lib.b200().rt_scene_inspect(C.byref(desc.c), C.byref(info)), "rt_scene_inspect")
    return info.as_dict()


def fp32_peak_tflops(device: int = -1) -> float:
    v = C.c_float(0.0)
    lib.check(lib.b200().rt_fp32_peak(device, C.byref(v)), "rt_fp32_peak")
    return float(v.value)
