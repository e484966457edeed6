"""Synthetic flattened scenes for the tests: many Translate(RotateY(primitive)) instances under one list -- the shape
the reference builds with Translate_new / RotateY_new (hittable.c:336-344, 369-389), in numbers its own eight scenes
never reach (two chains in the Cornell box, one in Book 2 final).  Only numpy + the ABI mirror: usable without a GPU.
"""
from __future__ import annotations

import numpy as np


def _bbox_of_points(pts):
    lo, hi = pts.min(axis=0), pts.max(axis=0)
    return np.array([lo[0], hi[0], lo[1], hi[1], lo[2], hi[2]], np.float32)


def _rot_y_corners(bbox, sn, cs):
    """Box of the 8 rotated corners, like RotateY_init (hittable.c:369-389)."""
    xs, ys, zs = bbox[0:2], bbox[2:4], bbox[4:6]
    pts = np.array([[cs * x + sn * z, y, -sn * x + cs * z] for x in xs for y in ys for z in zs], np.float32)
    return _bbox_of_points(pts)


def instanced_scene(rtc, base_desc, n_instances: int, seed: int = 7, with_quads: bool = True):
    """A world list of `n_instances` x Translate(RotateY(sphere | quad)) plus a few primitives in the world frame.
    Materials/textures are taken from `base_desc` (any built-in scene); returns a new SceneDesc."""
    abi = rtc.abi
    rng = np.random.default_rng(seed)
    n_mat = int(base_desc.c.n_materials)
    assert n_mat > 0
    nodes, spheres, quads, children = [], [], [], []

    def add_node(kind, a=0, b=0, flags=0, bbox=None, f=(0, 0, 0, 0)):
        n = np.zeros((), abi.NODE_DTYPE)
        n["kind"], n["a"], n["b"], n["flags"] = kind, a, b, flags
        n["bbox"] = bbox if bbox is not None else np.zeros(6, np.float32)
        n["f"] = np.asarray(f, np.float32)
        nodes.append(n)
        return len(nodes) - 1

    root = add_node(abi.NODE_LIST)  # children filled in below; DFS pre-order ids: the list first

    def add_sphere(center, radius):
        node_id = len(nodes)
        s = np.zeros((), abi.SPHERE_DTYPE)
        s["center"], s["radius"] = np.asarray(center, np.float32), np.float32(radius)
        s["material"], s["node"] = int(rng.integers(n_mat)), node_id
        spheres.append(s)
        c = np.asarray(center, np.float32)
        r = np.float32(radius)
        bbox = np.array([c[0] - r, c[0] + r, c[1] - r, c[1] + r, c[2] - r, c[2] + r], np.float32)
        return add_node(abi.NODE_SPHERE, a=len(spheres) - 1, bbox=bbox)

    def add_quad(Q, u, v):
        node_id = len(nodes)
        Q, u, v = (np.asarray(x, np.float32) for x in (Q, u, v))
        n = np.cross(u, v).astype(np.float32)
        area = np.float32(np.sqrt(np.float32(np.dot(n, n))))
        normal = (n * np.float32(1.0 / area)).astype(np.float32)  # vec3_div = mul by 1/s (vec3.c:18-19)
        q = np.zeros((), abi.QUAD_DTYPE)
        q["Q"], q["u"], q["v"], q["normal"] = Q, u, v, normal
        q["D"] = np.float32(np.dot(normal, Q))
        q["w"] = (n * np.float32(1.0 / np.float32(np.dot(n, n)))).astype(np.float32)
        q["area"], q["material"], q["node"] = area, int(rng.integers(n_mat)), node_id
        quads.append(q)
        pts = np.array([Q, Q + u, Q + v, Q + u + v], np.float32)
        bbox = _bbox_of_points(pts)
        for a in range(3):  # AABB_pad (hittable.c:24-37)
            if bbox[2 * a + 1] - bbox[2 * a] < 1e-4:
                bbox[2 * a] -= 1e-4
                bbox[2 * a + 1] += 1e-4
        return add_node(abi.NODE_QUAD, a=len(quads) - 1, bbox=bbox)

    top = []
    for k in range(n_instances):
        # Translate( RotateY( primitive ) ): pre-order ids translate < rotate < primitive
        t_id = add_node(abi.NODE_TRANSLATE)
        r_id = add_node(abi.NODE_ROTATE_Y)
        if with_quads and k % 3 == 2:
            p_id = add_quad(rng.uniform(-1, 1, 3), rng.uniform(0.5, 2, 3) * [1, 0.2, 0.1], rng.uniform(0.5, 2, 3) * [0.1, 1, 0.3])
        else:
            p_id = add_sphere(rng.uniform(-1, 1, 3), rng.uniform(0.3, 1.2))
        ang = np.float32(rng.uniform(-180, 180) * np.pi / 180.0)
        sn, cs = np.float32(np.sin(ang)), np.float32(np.cos(ang))
        off = rng.uniform(-15, 15, 3).astype(np.float32)
        nodes[r_id]["a"], nodes[r_id]["f"] = p_id, np.array([sn, cs, 0, 0], np.float32)
        nodes[r_id]["bbox"] = _rot_y_corners(nodes[p_id]["bbox"], sn, cs)
        nodes[t_id]["a"], nodes[t_id]["f"] = r_id, np.array([off[0], off[1], off[2], 0], np.float32)
        nodes[t_id]["bbox"] = nodes[r_id]["bbox"] + np.repeat(off, 2)
        top.append(t_id)
    for _ in range(6):  # world-frame primitives between them
        top.append(add_sphere(rng.uniform(-15, 15, 3), rng.uniform(0.5, 2.0)))
    if with_quads:
        top.append(add_quad([-20, -16, -20], [40, 0, 0], [0, 0, 40]))  # a floor
    children.extend(top)
    nodes[root]["a"], nodes[root]["b"] = 0, len(top)
    boxes = np.array([nodes[i]["bbox"] for i in top])
    nodes[root]["bbox"] = np.array([boxes[:, 0].min(), boxes[:, 1].max(), boxes[:, 2].min(), boxes[:, 3].max(),
                                    boxes[:, 4].min(), boxes[:, 5].max()], np.float32)

    out = abi.SceneDesc.from_bytes(base_desc.to_bytes())
    arrays = dict(out.arrays)
    arrays["nodes"] = np.array(nodes, abi.NODE_DTYPE)
    arrays["children"] = np.array(children, "<u4")
    arrays["spheres"] = np.array(spheres, abi.SPHERE_DTYPE) if spheres else np.zeros(0, abi.SPHERE_DTYPE)
    arrays["quads"] = np.array(quads, abi.QUAD_DTYPE) if quads else np.zeros(0, abi.QUAD_DTYPE)
    hdr = out.header
    hdr.n_nodes, hdr.n_children = len(nodes), len(children)
    hdr.n_spheres, hdr.n_quads = len(arrays["spheres"]), len(arrays["quads"])
    hdr.objects_root, hdr.lights_root = root, abi.RT_NONE
    out.arrays = arrays
    return abi.SceneDesc.from_bytes(out.to_bytes())
