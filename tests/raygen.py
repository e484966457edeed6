"""Deterministic ray batches for the hit-test parity checks: first-sample camera rays of a
scene plus secondary-like rays re-emitted from pseudo-random points, each with its own PCG32 state so that a
ConstantMedium's draw is reproducible per ray."""
from __future__ import annotations

import numpy as np


def _v(a):
    return np.array(list(a), dtype=np.float32)


def camera_rays(cam, n, seed=1):
    """n rays through pixel centres jittered like raytracing.c:104-122 (without the lens), float32 arithmetic."""
    rs = np.random.RandomState(seed)
    w, h = cam.img_width, cam.img_height
    i = rs.randint(0, w, size=n).astype(np.float32)
    j = rs.randint(0, h, size=n).astype(np.float32)
    jx = rs.uniform(-0.5, 0.5, size=n).astype(np.float32)
    jy = rs.uniform(-0.5, 0.5, size=n).astype(np.float32)
    p00, du, dv, eye = _v(cam.pixel00), _v(cam.pixel_delta_u), _v(cam.pixel_delta_v), _v(cam.origin)
    target = p00[None, :] + du[None, :] * (i + jx)[:, None] + dv[None, :] * (j + jy)[:, None]
    origins = np.repeat(eye[None, :], n, axis=0).astype(np.float32)
    dirs = (target - origins).astype(np.float32)
    return origins, dirs


def secondary_rays(desc, n, seed=2):
    """Rays starting inside the scene's bounding box (shrunk), uniformly random directions of random length."""
    rs = np.random.RandomState(seed)
    root = desc.nodes[desc.c.objects_root]
    bbox = np.array(root["bbox"], dtype=np.float64).reshape(3, 2)
    lo, hi = bbox[:, 0], bbox[:, 1]
    # the ground spheres make huge boxes: sample around the part of the scene where the objects are
    lo = np.maximum(lo, -600.0)
    hi = np.minimum(hi, 600.0)
    origins = rs.uniform(lo, hi, size=(n, 3)).astype(np.float32)
    d = rs.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    d *= rs.uniform(0.1, 10.0, size=(n, 1))
    return origins, d.astype(np.float32)


def surface_rays(desc, hits_p, hits_ok, n, seed=3):
    """Rays leaving previous hit points (what the integrator actually traces after the first bounce)."""
    rs = np.random.RandomState(seed)
    pts = hits_p[hits_ok]
    if len(pts) == 0:
        return secondary_rays(desc, n, seed)
    idx = rs.randint(0, len(pts), size=n)
    d = rs.normal(size=(n, 3))
    d /= np.linalg.norm(d, axis=1, keepdims=True)
    return pts[idx].astype(np.float32), d.astype(np.float32)


def rng_states(n, seed=4):
    rs = np.random.RandomState(seed)
    state = rs.randint(0, 2**63 - 1, size=n, dtype=np.int64).astype(np.uint64)
    inc = (rs.randint(0, 2**62, size=n, dtype=np.int64).astype(np.uint64) << np.uint64(1)) | np.uint64(1)
    return np.stack([state, inc], axis=1)


def edge_case_rays():
    """Axis-parallel and degenerate rays (d component = 0 -> 1/d = inf, 0*inf = NaN in the slab test), grazing and
    inside-sphere origins; scene independent."""
    o = [[0, 0, 0], [0, 1, 0], [13, 2, 3], [278, 278, -800], [278, 278, 278], [0, -1000, -1], [0, 0.2, 0],
         [4, 1, 0], [478, 278, -600], [0, 10, 0]]
    d = [[0, 0, -1], [1, 0, 0], [0, -1, 0], [0, 0, 1], [0, 1, 0], [1, 1, 0], [0, 0, 0], [-1, 0, 0], [0, 0, 1],
         [0, -1, 1e-30]]
    oo, dd = [], []
    for a in o:
        for b in d:
            oo.append(a)
            dd.append(b)
    return np.array(oo, np.float32), np.array(dd, np.float32)
