"""ctypes / numpy mirror of include/rtc_b200.h (the C ABI of librtc_b200.so).

Every layout here is checked against the header by tests/test_abi.py (sizeof of each struct through a tiny C
program), so a drifting field shows up as a failing CPU test rather than as garbage on the GPU.
"""
from __future__ import annotations

import ctypes as C
import io
from dataclasses import dataclass

import numpy as np

RT_ABI_VERSION = 1
RT_NONE = 0xFFFFFFFF

# rt_status
RT_OK, RT_ERR_INVALID, RT_ERR_NO_DEVICE, RT_ERR_CUDA, RT_ERR_UNSUPPORTED, RT_ERR_NOMEM = 0, -1, -2, -3, -4, -5
# rt_node_kind
NODE_LIST, NODE_SPHERE, NODE_QUAD, NODE_BVH, NODE_TRANSLATE, NODE_ROTATE_Y, NODE_MEDIUM = range(7)
# rt_material_tag (== the reference's MaterialType)
MAT_SURFACE_NORMAL, MAT_LAMBERTIAN, MAT_METAL, MAT_DIELECTRIC, MAT_DIFFUSE_LIGHT, MAT_ISOTROPIC = range(6)
# rt_texture_kind
TEX_SOLID, TEX_CHECKER, TEX_IMAGE, TEX_PERLIN = range(4)
# rt_kernel / rt_rng_mode / flags
KERNEL_EXACT, KERNEL_MEGA, KERNEL_WAVE = 0, 1, 2
RNG_PIXEL, RNG_SAMPLE = 0, 1
FLAG_DROP_NAN = 1

KERNELS = {"exact": KERNEL_EXACT, "mega": KERNEL_MEGA, "wave": KERNEL_WAVE}
RNG_MODES = {"pixel": RNG_PIXEL, "sample": RNG_SAMPLE}

NODE_DTYPE = np.dtype(
    [("kind", "<u4"), ("a", "<u4"), ("b", "<u4"), ("flags", "<u4"), ("bbox", "<f4", 6), ("f", "<f4", 4),
     ("reserved", "<u4", 2)]
)
SPHERE_DTYPE = np.dtype(
    [("center", "<f4", 3), ("radius", "<f4"), ("material", "<u4"), ("node", "<u4"), ("reserved", "<u4", 2)]
)
QUAD_DTYPE = np.dtype(
    [("Q", "<f4", 3), ("D", "<f4"), ("u", "<f4", 3), ("area", "<f4"), ("v", "<f4", 3), ("material", "<u4"),
     ("normal", "<f4", 3), ("node", "<u4"), ("w", "<f4", 3), ("reserved", "<u4")]
)
MATERIAL_DTYPE = np.dtype([("tag", "<u4"), ("texture", "<u4"), ("param", "<f4"), ("reserved", "<u4")])
TEXTURE_DTYPE = np.dtype(
    [("kind", "<u4"), ("a", "<u4"), ("b", "<u4"), ("scale", "<f4"), ("color", "<f4", 3), ("reserved", "<u4")]
)
PERLIN_DTYPE = np.dtype([("grad", "<f4", 768), ("perm_x", "<i4", 256), ("perm_y", "<i4", 256), ("perm_z", "<i4", 256)])
IMAGE_DTYPE = np.dtype([("width", "<i4"), ("height", "<i4"), ("offset", "<u8")])

assert NODE_DTYPE.itemsize == 64 and SPHERE_DTYPE.itemsize == 32 and QUAD_DTYPE.itemsize == 80
assert MATERIAL_DTYPE.itemsize == 16 and TEXTURE_DTYPE.itemsize == 32 and PERLIN_DTYPE.itemsize == 6144
assert IMAGE_DTYPE.itemsize == 16


class rt_scene_desc(C.Structure):
    _fields_ = [
        ("abi_version", C.c_uint32), ("n_nodes", C.c_uint32), ("n_children", C.c_uint32), ("n_spheres", C.c_uint32),
        ("n_quads", C.c_uint32), ("n_materials", C.c_uint32), ("n_textures", C.c_uint32), ("n_perlins", C.c_uint32),
        ("n_images", C.c_uint32), ("objects_root", C.c_uint32), ("lights_root", C.c_uint32), ("reserved", C.c_uint32),
        ("n_texel_bytes", C.c_uint64),
        ("nodes", C.c_void_p), ("children", C.c_void_p), ("spheres", C.c_void_p), ("quads", C.c_void_p),
        ("materials", C.c_void_p), ("textures", C.c_void_p), ("perlins", C.c_void_p), ("images", C.c_void_p),
        ("texels", C.c_void_p),
    ]


class rt_camera(C.Structure):
    _fields_ = [
        ("img_width", C.c_int32), ("img_height", C.c_int32), ("samples_per_pixel", C.c_int32), ("max_depth", C.c_int32),
        ("background", C.c_float * 3), ("lights_sampling_prob", C.c_float),
        ("origin", C.c_float * 3), ("dof_angle", C.c_float),
        ("pixel00", C.c_float * 3), ("reserved0", C.c_float),
        ("pixel_delta_u", C.c_float * 3), ("reserved1", C.c_float),
        ("pixel_delta_v", C.c_float * 3), ("reserved2", C.c_float),
        ("dof_disc_u", C.c_float * 3), ("reserved3", C.c_float),
        ("dof_disc_v", C.c_float * 3), ("reserved4", C.c_float),
    ]

    def copy(self) -> "rt_camera":
        other = rt_camera()
        C.memmove(C.byref(other), C.byref(self), C.sizeof(rt_camera))
        return other


class rt_opts(C.Structure):
    _fields_ = [
        ("kernel", C.c_uint32), ("rng_mode", C.c_uint32), ("sample_begin", C.c_int32), ("sample_count", C.c_int32),
        ("flags", C.c_uint32), ("reserved", C.c_uint32 * 3),
    ]


class rt_stats(C.Structure):
    _fields_ = [
        ("segments", C.c_uint64), ("light_pdf_evals", C.c_uint64), ("samples", C.c_uint64), ("nan_samples", C.c_uint64),
        ("kernel_ms", C.c_float), ("total_ms", C.c_float), ("launches", C.c_uint32), ("reserved", C.c_uint32),
    ]

    def as_dict(self) -> dict:
        return {name: getattr(self, name) for name, _ in self._fields_ if name != "reserved"}


class rt_scene_info(C.Structure):
    _fields_ = [
        ("bvh_nodes", C.c_uint32), ("bvh_depth", C.c_uint32), ("spheres", C.c_uint32), ("quads", C.c_uint32),
        ("instances", C.c_uint32), ("media", C.c_uint32), ("chain_ops", C.c_uint32), ("graph_frames", C.c_uint32),
        ("graph_rays", C.c_uint32), ("graph_recs", C.c_uint32), ("light_nesting", C.c_int32),
        ("mega_supported", C.c_uint32), ("device_bytes", C.c_uint64), ("pre_prims", C.c_uint32),
        ("wide_nodes", C.c_uint32),
    ]

    def as_dict(self) -> dict:
        return {name: getattr(self, name) for name, _ in self._fields_}


assert C.sizeof(rt_scene_info) == 64
assert C.sizeof(rt_scene_desc) == 128 and C.sizeof(rt_camera) == 128 and C.sizeof(rt_opts) == 32
assert C.sizeof(rt_stats) == 48

_BLOB_MAGIC = b"RTCSCN01"
_PARTS = [  # (field, count field, dtype) in the order the blob stores them
    ("nodes", "n_nodes", NODE_DTYPE),
    ("children", "n_children", np.dtype("<u4")),
    ("spheres", "n_spheres", SPHERE_DTYPE),
    ("quads", "n_quads", QUAD_DTYPE),
    ("materials", "n_materials", MATERIAL_DTYPE),
    ("textures", "n_textures", TEXTURE_DTYPE),
    ("perlins", "n_perlins", PERLIN_DTYPE),
    ("images", "n_images", IMAGE_DTYPE),
    ("texels", "n_texel_bytes", np.dtype("u1")),
]


@dataclass
class SceneDesc:
    """A flattened scene held in numpy arrays, with a ctypes view (`.c`) whose pointers stay valid while this
    object lives.  Built from the blob format of rtc_scene_desc_write() (flatten.c)."""

    header: rt_scene_desc
    arrays: dict

    @property
    def c(self) -> rt_scene_desc:
        return self.header

    def __getattr__(self, name):
        arrays = object.__getattribute__(self, "arrays")
        if name in arrays:
            return arrays[name]
        raise AttributeError(name)

    @staticmethod
    def from_bytes(blob: bytes) -> "SceneDesc":
        f = io.BytesIO(blob)
        if f.read(8) != _BLOB_MAGIC:
            raise ValueError("not an rtc scene blob")
        header = rt_scene_desc.from_buffer_copy(f.read(C.sizeof(rt_scene_desc)))
        if header.abi_version != RT_ABI_VERSION:
            raise ValueError(f"scene blob has ABI version {header.abi_version}")
        arrays = {}
        for field, count_field, dtype in _PARTS:
            n = getattr(header, count_field)
            raw = f.read(n * dtype.itemsize)
            if len(raw) != n * dtype.itemsize:
                raise ValueError("truncated scene blob")
            arr = np.frombuffer(raw, dtype=dtype).copy() if n else np.zeros(0, dtype=dtype)
            arrays[field] = arr
            setattr(header, field, arr.ctypes.data if n else None)
        return SceneDesc(header, arrays)

    @staticmethod
    def load(path) -> "SceneDesc":
        with open(path, "rb") as fh:
            return SceneDesc.from_bytes(fh.read())

    def to_bytes(self) -> bytes:
        header = rt_scene_desc.from_buffer_copy(bytes(self.header))
        for field, _, _ in _PARTS:
            setattr(header, field, None)
        return _BLOB_MAGIC + bytes(header) + b"".join(self.arrays[field].tobytes() for field, _, _ in _PARTS)

    def summary(self) -> dict:
        h = self.header
        return {k: getattr(h, k) for k in ("n_nodes", "n_spheres", "n_quads", "n_materials", "n_textures", "n_perlins",
                                            "n_images", "objects_root", "lights_root")}
