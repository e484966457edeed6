#!/usr/bin/env python
"""Megakernel throughput on the reference's eight scenes (`./main 0..7`), one GPU: the A/B table used when a kernel
change is measured.  One JSON line per scene: Mseg/s = segments / kernel time of the second (warm) render.

    python tools/scene_table.py [--width 600] [--spp 100] [--scenes 0,1,...] [--label text] [--reps 2]

RTC_BUILD_DIR selects another build of the library (see ray-tracing-c_b200/lib.py); the RT_MEGA_* switches apply.
"""
from __future__ import annotations

import argparse
import importlib
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
rtc = importlib.import_module("ray-tracing-c_b200")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--width", type=int, default=600)
    ap.add_argument("--spp", type=int, default=100)
    ap.add_argument("--scenes", default="0,1,2,3,4,5,6,7")
    ap.add_argument("--label", default="")
    ap.add_argument("--reps", type=int, default=2)
    ap.add_argument("--kernel", default="mega")
    args = ap.parse_args()
    row = {}
    for scene_id in [int(x) for x in args.scenes.split(",")]:
        sess = rtc.Session(scene_id, width=args.width, spp=args.spp)
        scene = rtc.Scene(sess.flatten())
        cam = sess.camera
        opts = rtc.make_opts(kernel=args.kernel, rng="sample")
        scene.render(cam, rtc.make_opts(kernel=args.kernel, rng="sample", sample_count=min(8, args.spp)),
                     want_radiance=False, want_rgb8=False)
        best = None
        for _ in range(args.reps):
            _, _, st = scene.render(cam, opts, want_radiance=False, want_rgb8=False)
            mseg = st["segments"] / st["kernel_ms"] * 1e-3
            best = mseg if best is None else max(best, mseg)
        row[scene_id] = round(best, 1)
        scene.close()
        sess.close()
    print(json.dumps({"label": args.label, "width": args.width, "spp": args.spp, "kernel": args.kernel,
                      "mseg_per_s": row}), flush=True)


if __name__ == "__main__":
    main()
