"""Developer diagnostic: prints the rays on which the mega kernel's hit batch differs from the oracle."""
from __future__ import annotations

import importlib
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
rtc = importlib.import_module("ray-tracing-c_b200")
import oracle_lib as O  # noqa: E402
import raygen  # noqa: E402
from conftest import load_golden  # noqa: E402


def main():
    for sid in [int(a) for a in sys.argv[1:]] or [0, 1, 5, 6]:
        sess = rtc.Session(sid, width=48, spp=4)
        desc = sess.flatten()
        scene = rtc.Scene(desc)
        g = load_golden(sid)
        o2, d2 = raygen.camera_rays(sess.camera, 200000, seed=201)
        o3, d3 = raygen.secondary_rays(desc, 200000, seed=202)
        origins = np.concatenate([g["rays_o"], o2, o3])
        dirs = np.concatenate([g["rays_d"], d2, d3])
        want = O.hit_batch(desc, origins, dirs)
        got = scene.hit_batch(origins, dirs, kernel="mega")
        bad = (want["node"] != got["node"]) | ~((want["t"].view(np.uint32) == got["t"].view(np.uint32)) |
                                                (np.isnan(want["t"]) & np.isnan(got["t"])))
        idx = np.nonzero(bad)[0]
        print(f"scene {sid}: {len(idx)} differing rays of {len(origins)}")
        for i in idx[:40]:
            print(f"   ray {i}: o={origins[i]} d={dirs[i]} oracle node {want['node'][i]} t {want['t'][i]!r} | "
                  f"mega node {got['node'][i]} t {got['t'][i]!r}")
        scene.close()

    # where does the end-to-end overhead go?
    sess = rtc.Session(1, width=1200, spp=8)
    for rep in range(3):
        t0 = time.perf_counter()
        desc = sess.flatten()
        t1 = time.perf_counter()
        scene = rtc.Scene(desc)
        t2 = time.perf_counter()
        _, _, st = scene.render(sess.camera, rtc.make_opts(kernel="mega", rng="sample"), want_radiance=False)
        t3 = time.perf_counter()
        scene.close()
        t4 = time.perf_counter()
        rgb, _, st2 = sess.camera_render(rtc.make_opts(kernel="mega", rng="sample"))
        t5 = time.perf_counter()
        print(f"flatten(py blob) {1e3 * (t1 - t0):.1f} ms, scene_create {1e3 * (t2 - t1):.1f} ms, render "
              f"{1e3 * (t3 - t2):.1f} ms (kernel {st['kernel_ms']:.1f}, total {st['total_ms']:.1f}), destroy "
              f"{1e3 * (t4 - t3):.1f} ms | camera_render {1e3 * (t5 - t4):.1f} ms (kernel {st2['kernel_ms']:.1f})")


if __name__ == "__main__":
    main()
