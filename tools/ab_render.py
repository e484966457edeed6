#!/usr/bin/env python
"""A/B timing of the render kernel for one build of the library (RTC_BUILD_DIR selects the build; the environment
carries the tuning knobs RT_MEGA_*).  Prints one JSON line per frame: Mseg/s from the CUDA-event kernel time.

    RTC_BUILD_DIR=ray-tracing-c_b200/_build_s0 python tools/ab_render.py --frames 1:1200:200,7:1000:100 --reps 3
"""
from __future__ import annotations

import argparse
import hashlib
import importlib
import json
import os
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
rtc = importlib.import_module("ray-tracing-c_b200")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--frames", default="1:1200:200,6:600:200,7:1000:100", help="scene:width:spp[:aspect],...")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--kernel", default="mega")
    ap.add_argument("--depth", type=int, default=50)
    ap.add_argument("--tag", default=os.environ.get("RTC_BUILD_DIR", "default"))
    args = ap.parse_args()
    for frame in args.frames.split(","):
        parts = frame.split(":")
        scene_id, width, spp = int(parts[0]), int(parts[1]), int(parts[2])
        aspect = float(parts[3]) if len(parts) > 3 else -1.0
        sess = rtc.Session(scene_id, width=width, spp=spp, depth=args.depth, aspect=aspect)
        scene = rtc.Scene(sess.flatten())
        opts = rtc.make_opts(kernel=args.kernel, rng="sample")
        rgb8, _, _ = scene.render(sess.camera, opts, want_radiance=False)
        sha = hashlib.sha256(rgb8.tobytes()).hexdigest()[:16]  # A/B builds must render the same frame
        best, seg = 1e30, 0
        for _ in range(args.reps):
            _, _, st = scene.render(sess.camera, opts, want_radiance=False)
            best, seg = min(best, st["kernel_ms"]), st["segments"]
        scene.close()
        knobs = {k: v for k, v in os.environ.items() if k.startswith("RT_MEGA")}
        print(json.dumps({"tag": args.tag, "knobs": knobs, "scene": scene_id, "frame": [sess.width, sess.height],
                          "spp": spp, "depth": args.depth, "sha": sha, "segments": seg, "kernel_ms": best, "mseg_per_s": seg / best * 1e-3}), flush=True)


if __name__ == "__main__":
    main()
