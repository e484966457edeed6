#!/usr/bin/env python
"""Per-scene throughput table for the configurations BASELINE.json names, on one GPU, next to the reference's
OpenMP build on the host cores.  Writes JSON lines (one per configuration x kernel) to stdout; the committed copies
live under profiles/.

    python tools/run_configs.py [--kernels mega,wave] [--quick] [--no-cpu]

cfg 1  ./main 0                500x281    100 spp          cfg 2  Book 1 final   1200x675  1000 spp
cfg 3  Cornell (mixture pdf)   600x600   1000 spp          cfg 4  Book 2 final   1000x1000 1000 spp
cfg 5  Book 2 final 4K         3840x2160 4096 spp -> rendered at 64 spp here (segments/s does not depend on spp; the
       full frame is 64x that work) unless --full5 is given
CPU: oracle/_ref/ref_main_fast (unmodified reference, -Ofast -fopenmp, all cores) at a bounded spp, segments from the
oracle's counters on the same frame at 1-2 spp.
"""
from __future__ import annotations

import argparse
import ctypes as C
import importlib
import json
import os
import subprocess
import sys
import tempfile
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))
rtc = importlib.import_module("ray-tracing-c_b200")

CONFIGS = [
    # name, scene, width, aspect, spp, cpu_spp
    ("cfg1 ./main 0 (defaults)", 0, 500, -1.0, 100, 100),
    ("cfg2 Book 1 final 1200x675x1000", 1, 1200, -1.0, 1000, 32),
    ("cfg3 Cornell box 600x600x1000", 6, 600, -1.0, 1000, 32),
    ("cfg4 Book 2 final 1000x1000x1000", 7, 1000, -1.0, 1000, 8),
    ("cfg5 Book 2 final 3840x2160x4096", 7, 3840, 16.0 / 9.0, 4096, 1),
]


def cpu_reference(scene, width, aspect, spp):
    """Seconds of the unmodified reference for the frame at `spp` (aspect overrides are not reachable through its
    CLI -- main.c:268 forces 1.0 for scene 7 -- so cfg 5 is timed on the 3840x3840 frame the CLI does produce and
    scaled by the pixel ratio)."""
    exe = ROOT / "oracle" / "_ref" / "ref_main_fast"
    if not exe.exists():
        return None
    cores = len(os.sched_getaffinity(0))
    env = dict(os.environ, OMP_NUM_THREADS=str(cores))
    with tempfile.TemporaryDirectory() as tmp:
        if scene in (3, 7):
            w, h = 1024, 512
            rgb = np.zeros((h, w, 3), np.uint8)
            rtc.lib.host().rtc_synthetic_image(w, h, rgb.ctypes.data_as(C.c_void_p))
            with open(Path(tmp) / "earthmap.jpg", "wb") as f:
                f.write(f"P6\n{w} {h}\n255\n".encode() + rgb.tobytes())
        t0 = time.perf_counter()
        subprocess.run([str(exe), str(scene), str(width), str(spp), "x"], cwd=tmp, env=env, check=True,
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        return time.perf_counter() - t0, cores


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--kernels", default="mega,wave")
    ap.add_argument("--quick", action="store_true", help="a tenth of the samples")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--full5", action="store_true")
    ap.add_argument("--only", default="")
    args = ap.parse_args()
    import oracle_lib as O

    for name, scene_id, width, aspect, spp, cpu_spp in CONFIGS:
        if args.only and args.only not in name:
            continue
        gpu_spp = spp
        if name.startswith("cfg5") and not args.full5:
            gpu_spp = 64
        if args.quick:
            gpu_spp = max(8, gpu_spp // 10)
        sess = rtc.Session(scene_id, width=width, spp=gpu_spp, aspect=aspect)
        desc, cam = sess.flatten(), sess.camera
        scene = rtc.Scene(desc)
        pixels = sess.width * sess.height

        # reference traversal work per segment (algorithmic FLOP model) from the oracle at 1 spp (small frame for 4K)
        osess = rtc.Session(scene_id, width=min(width, 600), spp=1, aspect=aspect)
        _, _, st_o = O.render(osess.flatten(), osess.camera, rtc.make_opts(kernel="exact", rng="pixel"), O.MATH_LIBM)
        work = O.last_work()
        seg_per_sample = st_o["segments"] / st_o["samples"]
        flop_per_seg = O.flop_per_segment(work)

        cpu = None
        if not args.no_cpu:
            ref_width = width
            scale = 1.0
            got = cpu_reference(scene_id, ref_width, aspect, cpu_spp)
            if got:
                sec, cores = got
                ref_pixels = pixels
                if aspect > 0:  # the reference CLI renders scene 7 square
                    ref_pixels = width * width
                segs = ref_pixels * cpu_spp * seg_per_sample
                cpu = {"mseg_per_s": segs / sec * 1e-6, "cores": cores, "spp": cpu_spp, "seconds": sec,
                       "full_frame_seconds_extrapolated": sec * (pixels / ref_pixels) * spp / cpu_spp}

        for kernel in args.kernels.split(","):
            opts = rtc.make_opts(kernel=kernel, rng="sample")
            scene.render(cam, rtc.make_opts(kernel=kernel, rng="sample", sample_count=min(8, gpu_spp)),
                         want_radiance=False)  # warm-up
            _, _, st = scene.render(cam, opts, want_radiance=False)
            mseg = st["segments"] / st["kernel_ms"] * 1e-3
            line = {
                "config": name, "scene": scene_id, "image": [sess.width, sess.height], "spp_rendered": gpu_spp,
                "spp_config": spp, "kernel": kernel, "segments": st["segments"], "kernel_ms": st["kernel_ms"],
                "total_ms": st["total_ms"], "mseg_per_s": mseg, "seg_per_sample": st["segments"] / st["samples"],
                # kernel time scales with spp; upload, tonemap and the copies back are per frame
                "full_frame_seconds": (st["kernel_ms"] * spp / gpu_spp + (st["total_ms"] - st["kernel_ms"])) * 1e-3,
                "flop_per_segment_reference_traversal": flop_per_seg,
                "algorithmic_tflops": mseg * 1e6 * flop_per_seg * 1e-12, "nan_samples": st["nan_samples"],
                "cpu_reference": cpu,
                "speedup_vs_cpu": (mseg / cpu["mseg_per_s"]) if cpu else None,
            }
            print(json.dumps(line), flush=True)
        scene.close()


if __name__ == "__main__":
    main()
