#!/usr/bin/env python
"""Diagnostic for tests/test_gpu_mega.py::test_hit_parity_with_media: where do the megakernel's medium decisions differ
from the oracle's?  Prints counts by category and the first few offending rays."""
import importlib, sys
from pathlib import Path
import numpy as np
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
rtc = importlib.import_module("ray-tracing-c_b200")
import oracle_lib as O, raygen
from conftest import load_golden

MUL = np.uint64(6364136223846793005)
def steps(before, after, inc):
    s = before.copy(); out = np.full(len(s), -1)
    with np.errstate(over="ignore"):
        for k in range(5):
            out[(out < 0) & (s == after)] = k
            s = s * MUL + inc
    return out

sess = rtc.Session(7, width=48, spp=4); desc = sess.flatten(); scene = rtc.Scene(desc)
g = load_golden(7)
o2, d2 = raygen.camera_rays(sess.camera, 150_000, seed=201); o3, d3 = raygen.secondary_rays(desc, 150_000, seed=202)
origins = np.concatenate([g["rays_o"], o2, o3]); dirs = np.concatenate([g["rays_d"], d2, d3])
real = (dirs != 0).any(axis=1); origins, dirs = origins[real], dirs[real]
rng = raygen.rng_states(len(origins), seed=203)
want = O.hit_batch(desc, origins, dirs, rng=rng, math=O.MATH_PORTABLE)
got = scene.hit_batch(origins, dirs, rng=rng, kernel="mega")
kinds = desc.nodes["kind"]; med_ids = np.flatnonzero(kinds == rtc.abi.NODE_MEDIUM)
print("media node ids", med_ids, "n_nodes", len(kinds))
dw = steps(rng[:, 0], want["rng"][:, 0], rng[:, 1]); dg = steps(rng[:, 0], got["rng"][:, 0], rng[:, 1])
bad_rng = dw != dg; bad_node = want["node"] != got["node"]
bad_t = want["t"].view(np.uint32) != got["t"].view(np.uint32)
print("rays", len(origins), "draw-count mismatches", int(bad_rng.sum()), "node mismatches", int(bad_node.sum()), "t mismatches", int((bad_t & ~bad_node).sum()))
print("draws want histogram", np.bincount(dw + 1), "got", np.bincount(dg + 1))
import collections
cat = collections.Counter()
for i in np.flatnonzero(bad_rng | bad_node | bad_t):
    wn, gn = int(want["node"][i]), int(got["node"][i])
    def cls(n):
        if n == 0xFFFFFFFF: return "miss"
        if kinds[n] == rtc.abi.NODE_MEDIUM: return f"medium{list(med_ids).index(n)}"
        return "early" if n < med_ids.min() else "late"
    cat[(cls(wn), cls(gn), int(dw[i]), int(dg[i]))] += 1
for k, v in sorted(cat.items(), key=lambda kv: -kv[1]): print(v, "want/got class, draws want/got:", k)
for i in np.flatnonzero(bad_rng | bad_node | bad_t)[:12]:
    print(i, "o", origins[i], "d", dirs[i], "want", want["node"][i], float(want["t"][i]), dw[i], "got", got["node"][i], float(got["t"][i]), dg[i])
