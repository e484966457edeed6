"""Developer diagnostic (not a pytest file): runs every scene through the exact and mega kernels on cuda:0 and
prints how they compare with the CPU oracle.  `python tools/dev/dev_check.py [scene ...]`"""
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


def ulp_diff(a, b):
    ia = a.view(np.int32).astype(np.int64)
    ib = b.view(np.int32).astype(np.int64)
    ia = np.where(ia < 0, -(ia & 0x7FFFFFFF), ia)
    ib = np.where(ib < 0, -(ib & 0x7FFFFFFF), ib)
    return np.abs(ia - ib)


def main():
    scenes = [int(a) for a in sys.argv[1:]] or list(range(8))
    for sid in scenes:
        w = 96 if sid != 7 else 64
        spp = 8
        sess = rtc.Session(sid, width=w, spp=spp)
        desc = sess.flatten()
        cam = sess.camera
        info = rtc.inspect(desc)
        scene = rtc.Scene(desc)
        print(f"== scene {sid}: {cam.img_width}x{cam.img_height} spp {spp}  bvh nodes {info['bvh_nodes']} "
              f"depth {info['bvh_depth']}")

        # ---- hit batches
        n = 20000
        o1, d1 = raygen.camera_rays(cam, n)
        o2, d2 = raygen.secondary_rays(desc, n)
        oe, de = raygen.edge_case_rays()
        origins = np.concatenate([o1, o2, oe])
        dirs = np.concatenate([d1, d2, de])
        rng = raygen.rng_states(len(origins))
        ref = O.hit_batch(desc, origins, dirs, rng=rng)
        for kern in ("exact", "mega"):
            got = scene.hit_batch(origins, dirs, rng=rng, kernel=kern)
            hit_r, hit_g = ref["node"] != 0xFFFFFFFF, got["node"] != 0xFFFFFFFF
            both = hit_r & hit_g
            ulps = ulp_diff(ref["t"][both], got["t"][both])
            print(f"   hits[{kern}]: rays {len(origins)} ref-hits {hit_r.sum()} hit-mismatch {(hit_r != hit_g).sum()} "
                  f"id-mismatch {(ref['node'][both] != got['node'][both]).sum()} max-ulp(t) "
                  f"{ulps.max() if len(ulps) else 0} t-bitdiff {(ulps != 0).sum()} "
                  f"p-bitdiff {(ref['p'][both].view(np.uint32) != got['p'][both].view(np.uint32)).any(axis=1).sum()} "
                  f"n-bitdiff {(ref['normal'][both].view(np.uint32) != got['normal'][both].view(np.uint32)).any(axis=1).sum()} "
                  f"uv-maxdiff {np.nanmax(np.abs(ref['uv'][both] - got['uv'][both])) if both.any() else 0:.3g} "
                  f"rng-diff {(ref['rng'][:, 0] != got['rng'][:, 0]).sum()}")

        # ---- renders
        for rng_mode in ("pixel", "sample"):
            opts = rtc.make_opts(kernel="exact", rng=rng_mode)
            t0 = time.time()
            _, rad_o, st_o = O.render(desc, cam, opts, O.MATH_PORTABLE)
            t_or = time.time() - t0
            _, rad_g, st_g = scene.render(cam, opts)
            same = rad_o.view(np.uint32) == rad_g.view(np.uint32)
            eq = (rad_o == rad_g) | (np.isnan(rad_o) & np.isnan(rad_g))
            print(f"   render[exact,{rng_mode}]: bit-identical {same.all()} equal {eq.all()} differing values "
                  f"{(~eq).sum()}/{eq.size} seg oracle {st_o['segments']} gpu {st_g['segments']} "
                  f"kernel {st_g['kernel_ms']:.1f} ms (oracle {t_or * 1e3:.0f} ms)")
        opts = rtc.make_opts(kernel="mega", rng="sample")
        _, rad_m, st_m = scene.render(cam, opts)
        _, rad_o, st_o = O.render(desc, cam, rtc.make_opts(kernel="exact", rng="sample"), O.MATH_PORTABLE)
        a, b = rad_m / spp, rad_o / spp
        ok = np.isfinite(a) & np.isfinite(b)
        rmse = float(np.sqrt(np.mean((np.clip(a[ok], 0, 1) - np.clip(b[ok], 0, 1)) ** 2)))
        print(f"   render[mega,sample]: rmse vs oracle {rmse:.4f} mean {a[ok].mean():.4f} vs {b[ok].mean():.4f} "
              f"segments {st_m['segments']} vs {st_o['segments']} nan {st_m['nan_samples']} vs {st_o['nan_samples']} "
              f"kernel {st_m['kernel_ms']:.2f} ms -> {st_m['segments'] / max(st_m['kernel_ms'], 1e-6) * 1e-3:.1f} Mseg/s")
        scene.close()

    # throughput at the headline config (short): scene 1, 1200x675
    if not sys.argv[1:]:
        sess = rtc.Session(1, width=1200, spp=64)
        scene = rtc.Scene(sess.flatten())
        for rep in range(3):
            _, _, st = scene.render(sess.camera, rtc.make_opts(kernel="mega", rng="sample"), want_radiance=False)
            print(f"scene 1 1200x675x64: kernel {st['kernel_ms']:.1f} ms total {st['total_ms']:.1f} ms "
                  f"{st['segments'] / st['kernel_ms'] * 1e-3:.1f} Mseg/s")
        print("fp32 peak TFLOP/s:", rtc.fp32_peak_tflops())


if __name__ == "__main__":
    main()
