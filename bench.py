#!/usr/bin/env python
"""bench.py -- throughput of the per-pixel render loop on B200s, beside the reference's OpenMP build.

    python bench.py --gpus N --steps K --warmup W            this backend (librtc_b200.so through the C ABI)
    python bench.py --impl reference --gpus N --steps K ...  the UNMODIFIED reference (oracle/_ref/ref_main_fast,
                                                             -Ofast -fopenmp, all host cores) on the same config

Workload (default = BASELINE.json configs[1], the configuration the metric is quoted on): the Book 1 final scene
(`./main 1 1200 1000 x`): 486 spheres, 1200x675, 1000 samples per pixel, depth 50.  `--config 1..5` selects any of
BASELINE.json's five configurations (5 = Book 2 final at 3840x2160, 4096 spp: the aspect override the reference's
CLI cannot express, src/main.c:268).  One STEP = one whole render of that frame.  For N > 1 (launched by torchrun,
one rank per GPU) the samples of every pixel are split into N contiguous ranges (strong scaling: the frame is
fixed), each rank renders its range into a float radiance buffer, one NCCL sum-reduce brings them to rank 0, which
tonemaps.  After the timed regions rank 0 renders the WHOLE sample range on its own GPU (untimed) and compares it
with the N-rank image (`image_check`); a difference of more than one 8-bit level fails the run.

Printed line (rank 0): metric = path segments per second (the integrator's closest-hit queries, raytracing.c:45),
`value` with the scene resident in HBM, `e2e` through the reference-facing call with HOST buffers
(Camera_render: World graph in host memory -> 8-bit image in host memory, flatten + upload + download inside the
timed region), `roofline` of the megakernel, `cpu_baseline` = the reference binary timed on this box's cores.
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
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

UNIT = "Mseg/s"
DEPTH = 50
# BASELINE.json configs[0..4] -> --config 1..5.  ref_spp: samples per pixel of one step of the CPU reference arm
# (a bounded sample of the same frame: cost is linear in spp, raytracing.c:100), sized for 5-20 s on 16-32 cores.
CONFIGS = {
    1: dict(scene=0, width=500, spp=100, aspect=-1.0, ref_spp=100, title="./main 0 (reference defaults)"),
    2: dict(scene=1, width=1200, spp=1000, aspect=-1.0, ref_spp=64, title="Book 1 final scene"),
    3: dict(scene=6, width=600, spp=1000, aspect=-1.0, ref_spp=64, title="Cornell box (mixture pdf)"),
    4: dict(scene=7, width=1000, spp=1000, aspect=-1.0, ref_spp=16, title="Book 2 final scene"),
    5: dict(scene=7, width=3840, spp=4096, aspect=16.0 / 9.0, ref_spp=2, title="Book 2 final scene"),
}
REF_FAST = ROOT / "oracle" / "_ref" / "ref_main_fast"
REF_TOOL_FAST = ROOT / "oracle" / "_ref" / "ref_tool_fast"
# ncu counters of the bench's own launches (ncu --set full, one capture per configuration), written by
# `profiles/summarize_ncu.py json`; bench.py only quotes them for the configuration they were captured on
NCU_SUMMARY = ROOT / "profiles" / "ncu_summary.json"
FALLBACK_FLOP_PER_SEGMENT = {0: 196.0, 1: 1434.0, 6: 828.0, 7: 1277.0}  # SURVEY.md 8(d); recomputed live when possible


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--kernel", default="mega", choices=["mega", "exact", "wave"])
    ap.add_argument("--config", type=int, default=2, choices=sorted(CONFIGS),
                    help="BASELINE.json configuration 1..5 (default 2: the one the metric is quoted on)")
    # overrides for development only; a line produced with these says so in config.workload
    ap.add_argument("--scene", type=int, default=None)
    ap.add_argument("--width", type=int, default=None)
    ap.add_argument("--spp", type=int, default=None)
    ap.add_argument("--aspect", type=float, default=None, help="aspect ratio override (config 5 uses 16/9)")
    ap.add_argument("--ref-spp", type=int, default=0, help="samples per pixel of one reference/CPU step (0 = auto)")
    ap.add_argument("--warmup-spp", type=int, default=0,
                    help="samples per pixel of the warm-up steps (0 = the full frame; for the minute-long config 5)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-image-check", action="store_true")
    ap.add_argument("--no-dropin", action="store_true", help="skip the RT_GPUS=N Camera_render measurement at N > 1")
    args = ap.parse_args()
    cfg = CONFIGS[args.config]
    args.overridden = any(v is not None for v in (args.scene, args.width, args.spp, args.aspect))
    args.scene = cfg["scene"] if args.scene is None else args.scene
    args.width = cfg["width"] if args.width is None else args.width
    args.spp = cfg["spp"] if args.spp is None else args.spp
    args.aspect = cfg["aspect"] if args.aspect is None else args.aspect
    args.title = cfg["title"] if args.scene == cfg["scene"] else f"scene {args.scene}"
    return args


def open_session(args, spp=None):
    rtc = importlib.import_module("ray-tracing-c_b200")
    return rtc.Session(args.scene, width=args.width, spp=spp or args.spp, depth=DEPTH, aspect=args.aspect)


def frame_size(args):
    sess = open_session(args, spp=1)
    return sess.width, sess.height


def workload_name(args):
    w, h = frame_size(args)
    tag = f"config {args.config}" + (" (overridden: development frame)" if args.overridden else "")
    return f"{tag}: {args.title}, {w}x{h}, {args.spp} spp, depth {DEPTH}"


def metric_name(args):
    w, h = frame_size(args)
    return f"path segments per second (Mrays/s), {args.title} {w}x{h}, {args.spp} spp, depth {DEPTH}"


# ---------------------------------------------------------------------- clocks --------------------------------------

class ClockSampler:
    """nvidia-smi polled every 200 ms while the timed region runs (the recipe's clocks line)."""

    QUERY = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu_index = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.QUERY}", "--format=csv,noheader,nounits", "-lms", "200", "-i",
                 str(self.gpu_index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, sm_max, power, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for line in self.lines:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 9:
                continue
            try:
                sm.append(float(parts[1]))
                sm_max.append(float(parts[2]))
                power.append(float(parts[3]))
            except ValueError:
                continue
            for name, flag in zip(names, parts[5:9]):
                if flag.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(sm_max)), "power_w_max": float(max(power)),
                "samples": len(sm), "reasons": sorted(reasons)}


# ---------------------------------------------------------------------- the reference on the host cores --------------

def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def oracle_segments_per_sample(args):
    """Segments per camera sample and algorithmic FLOP per segment of the REFERENCE's traversal, from the oracle's
    counters (oracle/rtc_oracle.c, LIBM mode == the compiled reference) on the frame at 2 spp."""
    sys.path.insert(0, str(ROOT / "tests"))
    import oracle_lib as O

    rtc = O.rtc
    w, h = frame_size(args)
    # a frame of at most ~0.6 Mpixel at 2 spp (the 8-Mpixel frame of config 5 is counted on a quarter-size frame of
    # the same view: segments per sample do not depend on the pixel pitch)
    scale = max(1, int(np.ceil(np.sqrt(w * h / 6.0e5))))
    sess = rtc.Session(args.scene, width=max(16, args.width // scale), spp=2, depth=DEPTH, aspect=args.aspect)
    desc = sess.flatten()
    _, _, st = O.render(desc, sess.camera, rtc.make_opts(kernel="exact", rng="pixel"), O.MATH_LIBM)
    work = O.last_work()
    return st["segments"] / st["samples"], O.flop_per_segment(work), work


def run_reference_once(args, spp: int, threads: int):
    """(seconds inside Camera_render, wall seconds of the process) of the unmodified reference at `spp`.

    oracle/_ref/ref_tool_fast = the reference's sources with its own flags (-Ofast -fopenmp) behind oracle/ref_tool.c:
    it times Camera_render alone with a monotonic clock -- the scope of the reference's own timer (src/main.c:334-338:
    scene build, process start and the TIFF write are outside) -- and takes the aspect override of config 5.  Without
    it: wall clock around the reference's CLI (ref_main_fast), fixed costs inside."""
    env = dict(os.environ, OMP_NUM_THREADS=str(threads))
    with tempfile.TemporaryDirectory() as tmp:
        if REF_TOOL_FAST.exists():  # (writes the synthetic earthmap.jpg itself for scenes 3 and 7)
            t0 = time.perf_counter()
            out = subprocess.run([str(REF_TOOL_FAST), "render", str(args.scene), str(args.width), str(spp), str(DEPTH),
                                  f"{max(args.aspect, 0.0):.9g}", "out.rgb8", "out.cam"], cwd=tmp, env=env, check=True,
                                 stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            wall = time.perf_counter() - t0
            inner = [float(l.split("=")[1]) for l in out.stdout.splitlines() if l.startswith("render_seconds=")]
            return (inner[0] if inner else wall), wall
        if args.scene in (3, 7):  # the asset the reference loads from its cwd (absent from its repo): synthetic image
            rtc = importlib.import_module("ray-tracing-c_b200")
            w, h = 1024, 512
            rgb = np.zeros((h, w, 3), np.uint8)
            rtc.lib.host().rtc_synthetic_image(w, h, rgb.ctypes.data_as(C.c_void_p))
            with open(Path(tmp) / "earthmap.jpg", "wb") as f:
                f.write(f"P6\n{w} {h}\n255\n".encode() + rgb.tobytes())
        t0 = time.perf_counter()
        subprocess.run([str(REF_FAST), str(args.scene), str(args.width), str(spp), "x"], cwd=tmp, env=env, check=True,
                       stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
        wall = time.perf_counter() - t0
        return wall, wall


def run_port_once(args, spp: int) -> float:
    """Fallback when the reference binary is absent: the oracle port (OpenMP over rows)."""
    sys.path.insert(0, str(ROOT / "tests"))
    import oracle_lib as O

    rtc = O.rtc
    sess = open_session(args, spp=spp)
    desc = sess.flatten()
    t0 = time.perf_counter()
    O.render(desc, sess.camera, rtc.make_opts(kernel="exact", rng="pixel"), O.MATH_LIBM)
    sec = time.perf_counter() - t0
    return sec, sec


def pixels_of(args) -> int:
    w, h = frame_size(args)
    return w * h


def cpu_reference_arm(args, steps: int, warmup: int, spp: int):
    """Times `steps` bounded samples of the workload on the host cores.  Returns (Mseg/s, ms per step, descriptor)."""
    cores = host_cores()
    seg_per_sample, flop_per_seg, _ = oracle_segments_per_sample(args)
    have_ref = REF_TOOL_FAST.exists() or (REF_FAST.exists() and args.aspect <= 0)
    kind = "reference" if have_ref else "port"
    run = (lambda: run_reference_once(args, spp, cores)) if kind == "reference" else (lambda: run_port_once(args, spp))
    for _ in range(warmup):
        run()
    times = [run() for _ in range(max(steps, 1))]
    w, h = frame_size(args)
    segments = w * h * spp * seg_per_sample
    sec = float(np.mean([t[0] for t in times]))
    wall = float(np.mean([t[1] for t in times]))
    value = segments / sec * 1e-6
    if kind == "reference" and REF_TOOL_FAST.exists():
        how = ("oracle/_ref/ref_tool_fast (the unmodified reference sources, gcc -Ofast -fopenmp, behind oracle/ref_tool.c)"
               f" `render {args.scene} {args.width} {spp} {DEPTH} {max(args.aspect, 0.0):.4g}`; timed: Camera_render alone,"
               " monotonic clock (the scope of the reference's own timer, src/main.c:334-338)")
    elif kind == "reference":
        how = (f"oracle/_ref/ref_main_fast (unmodified reference, gcc -Ofast -fopenmp) `main {args.scene} {args.width} "
               f"{spp} x`; timed: wall clock of the process (scene build and TIFF write inside)")
    else:
        how = "oracle port (liboracle.so, OpenMP over rows); timed: the render call"
    desc = {
        "value": value, "unit": UNIT, "cores": cores, "kind": kind,
        "sample": (f"{how}; the full {w}x{h} frame at {spp} of {args.spp} spp, {sec:.2f} s per run, mean of "
                   f"{len(times)}; segments = pixels*spp*{seg_per_sample:.3f} seg/sample counted by the oracle on the same view"),
        "seconds_per_run": sec, "process_wall_seconds": wall, "spp": spp, "seg_per_sample": seg_per_sample,
        "full_frame_seconds_extrapolated": sec * args.spp / spp,
    }
    return value, sec * 1e3, desc, flop_per_seg


# ---------------------------------------------------------------------- this backend --------------------------------

def ncu_summary(args, world, variant_threads):
    """The committed ncu counters of this configuration's launch (profiles/ncu_summary.json), or None.  They are only
    quoted when the capture is of the SAME configuration (frame, spp) -- per-rank figures are scaled by the rank's
    share of the segments, everything else is reported as null."""
    try:
        table = json.loads(NCU_SUMMARY.read_text())
    except (OSError, ValueError):
        return None
    entry = table.get(f"config{args.config}")
    if entry is None or args.overridden:
        return None
    return entry


def image_diff(a: np.ndarray, b: np.ndarray) -> dict:
    import hashlib

    d = a.astype(np.int16) - b.astype(np.int16)
    return {"max_abs_diff_u8": int(np.abs(d).max()) if d.size else 0,
            "rmse_u8": float(np.sqrt(np.mean(d.astype(np.float64) ** 2))) if d.size else 0.0,
            "differing_bytes": int(np.count_nonzero(d)), "bytes": int(d.size),
            "sha256": hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()}


def b200_arm(args):
    import torch
    import torch.distributed as td

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("bench.py --gpus N>1 must be launched by torchrun (one rank per GPU)")
        args.gpus = world
    torch.cuda.set_device(local_rank)
    cpu_group = None
    if world > 1:
        td.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
        cpu_group = td.new_group(backend="gloo")  # host-side waits (a NCCL barrier would spin ON the waiting GPUs)

    rtc = importlib.import_module("ray-tracing-c_b200")
    dist = importlib.import_module("ray-tracing-c_b200.dist")
    sess = open_session(args)
    desc, cam = sess.flatten(), sess.camera
    info = rtc.inspect(desc)
    renderer = dist.DistRenderer(desc, cam, device=local_rank)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2

    def barrier():
        if world > 1:
            td.barrier()
        torch.cuda.synchronize()

    # ---- resident-scene throughput (`value`)
    warm_spp = args.warmup_spp if 0 < args.warmup_spp < args.spp else 0
    for _ in range(args.warmup):
        renderer.enqueue(kernel=args.kernel, spp=warm_spp or None)
        torch.cuda.synchronize()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    step_ms, kernel_ms, reduce_ms, seg_local, launches = [], [], [], 0, 0
    for _ in range(args.steps):
        flush.zero_()  # L2 flush between timed iterations (untimed)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        renderer.enqueue(kernel=args.kernel)
        e1.record()
        st = renderer.stats()  # synchronises; kernel_ms comes from CUDA events around the kernel on its own stream
        step_ms.append(e0.elapsed_time(e1))
        kernel_ms.append(st["kernel_ms"])
        reduce_ms.append(st.get("reduce_ms", 0.0))
        seg_local = st["segments"]
        launches += st["launches"] + (1 if rank == 0 else 0)  # + the tonemap on rank 0
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    image_dist = renderer.host_rgb8.numpy().copy() if rank == 0 else None  # the N-rank image of the last timed step
    total_ms = torch.tensor([sum(step_ms)], dtype=torch.float64, device="cuda")
    if world > 1:
        td.all_reduce(total_ms, op=td.ReduceOp.MAX)
    ms_per_step = float(total_ms.item()) / args.steps
    segments = renderer.total_segments(seg_local)
    value = segments / ms_per_step * 1e-3
    red_t = torch.tensor([float(np.mean(reduce_ms))], dtype=torch.float64, device="cuda")
    if world > 1:
        td.all_reduce(red_t, op=td.ReduceOp.MAX)

    # ---- end to end through the reference-facing call, host buffers in and out (`e2e`)
    e2e_steps = min(args.steps, 3)
    h, w = cam.img_height, cam.img_width
    barrier()
    e2e_ms = []
    e2e_warm = max(1, min(args.warmup, 3))
    sess_warm = open_session(args, spp=warm_spp) if warm_spp else sess
    rgb8_e2e = None
    for it in range(-e2e_warm, e2e_steps):  # negative iterations: untimed warm-up (first-use allocations of the pools)
        barrier()
        t0 = time.perf_counter()
        if world == 1:
            # Camera_render itself: World/Camera structs in host memory -> uint8 image in host memory
            rgb8_e2e, _, st_e = (sess if it >= 0 else sess_warm).camera_render(rtc.make_opts(kernel=args.kernel, rng="sample"))
        else:
            # from the host description: the uploaded scene comes from the library's cache, as in Camera_render
            r2 = dist.DistRenderer(desc, cam, device=local_rank, cached=True)
            rgb8_e2e, st_e = r2.render(kernel=args.kernel, spp=None if it >= 0 else (warm_spp or None))
            r2.close()
        barrier()
        if it >= 0:
            e2e_ms.append((time.perf_counter() - t0) * 1e3)
    e2e_t = torch.tensor([float(np.mean(e2e_ms))], dtype=torch.float64, device="cuda")
    if world > 1:
        td.all_reduce(e2e_t, op=td.ReduceOp.MAX)
    e2e_value = segments / float(e2e_t.item()) * 1e-3

    # ---- is the image right?  N > 1: rank 0 renders the WHOLE sample range on its own GPU (untimed) and compares it
    # with the NCCL-reduced image of the N ranks.  N = 1: the resident-scene image against Camera_render's.
    image_check = None
    if rank == 0 and not args.no_image_check:
        if world > 1:
            whole, whole_rad = renderer.render_whole(kernel=args.kernel)
            image_check = image_diff(image_dist, whole)
            red = renderer.radiance.cpu().numpy()
            scale = np.maximum(np.abs(whole_rad), 1e-3 * args.spp)
            rel = np.abs(red - whole_rad) / scale
            finite = np.isfinite(rel)  # a NaN sample poisons its pixel's sum on both sides (as in the reference)
            image_check["radiance_max_rel_diff"] = float(rel[finite].max()) if finite.any() else None
            image_check["non_finite_values"] = [int((~np.isfinite(red)).sum()), int((~np.isfinite(whole_rad)).sum())]
            image_check["against"] = f"all {args.spp} spp rendered by rank 0's GPU alone (same per-(pixel,sample) streams)"
        else:
            image_check = image_diff(image_dist, rgb8_e2e)
            image_check["against"] = "Camera_render (host buffers) vs the resident-scene path"
    if cpu_group is not None:
        td.barrier(group=cpu_group)

    # ---- the literal drop-in at N > 1: ONE process, Camera_render with RT_GPUS=N (rt_render_multi: in-process NCCL).
    # Rank 0 alone runs it; the other ranks idle at a host-side barrier (their GPUs are free).
    dropin = None
    if world > 1 and not args.no_dropin:
        if rank == 0:
            try:
                times = []
                for it in range(-2, min(args.steps, 3)):
                    t0 = time.perf_counter()
                    rgb8_d, _, st_d = (sess if it >= 0 else sess_warm).camera_render(
                        rtc.make_opts(kernel=args.kernel, rng="sample"), n_gpus=world)
                    if it >= 0:
                        times.append((time.perf_counter() - t0) * 1e3)
                ms = float(np.mean(times))
                dropin = {"value": segments / ms * 1e-3, "unit": UNIT, "ms_per_step": ms,
                          "ms_steps": [round(v, 2) for v in times], "kernel_ms_max": st_d.get("kernel_ms"),
                          "api": f"Camera_render with RT_GPUS={world} (rt_render_multi: one process, in-process NCCL reduce)",
                          "image_vs_ranks": image_diff(rgb8_d, image_dist) if image_dist is not None else None}
            except Exception as exc:  # a failure here is reported, it does not void the torchrun numbers
                dropin = {"error": str(exc)[:300]}
        td.barrier(group=cpu_group)

    if rank != 0:
        renderer.close()
        if world > 1:
            td.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (mega_render_kernel): FP32 issue, not HBM and not tensor
    fp32_peak = rtc.fp32_peak_tflops(local_rank)
    cpu = None
    flop_per_seg = FALLBACK_FLOP_PER_SEGMENT.get(args.scene, 1434.0)
    flop_source = "SURVEY.md 8(d) constant"
    if not args.no_cpu_baseline:
        try:
            if world == 1:
                ref_spp = args.ref_spp or CONFIGS[args.config]["ref_spp"]
                _, _, cpu, flop_per_seg = cpu_reference_arm(args, steps=1, warmup=0, spp=min(ref_spp, args.spp))
            else:
                _, flop_per_seg, _ = oracle_segments_per_sample(args)  # counters only; the CPU timing is N = 1's job
            flop_source = "oracle counters on this view (tests/oracle_lib.py: flop_per_segment)"
        except Exception as exc:
            flop_source = f"SURVEY.md 8(d) constant (oracle unavailable: {str(exc)[:80]})"
    k_ms = float(np.mean(kernel_ms))
    achieved_tflops = seg_local * flop_per_seg / (k_ms * 1e-3) * 1e-12
    n_values = h * w * 3
    peaks = {}
    try:
        peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text())
    except (OSError, ValueError):
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    # compulsory HBM bytes of one launch: the scene read once and the summed frame written once (the per-chunk partial
    # sums live in L2-sized buffers and are an implementation choice)
    hbm_bytes = n_values * 4 + int(info["device_bytes"])
    ncu = ncu_summary(args, world, None)
    share = None
    if ncu and ncu.get("segments"):
        share = seg_local / float(ncu["segments"])  # this rank's launch against the captured one (N = 1, full frame)
    roofline = {
        "bound": "fp32", "achieved": achieved_tflops, "peak": fp32_peak, "unit": "TFLOP/s",
        "frac": achieved_tflops / fp32_peak if fp32_peak > 0 else None,
        "frac_is": "ALGORITHMIC credit: segments x FLOP of the REFERENCE's own traversal per segment (its 43 box tests "
                   "per segment on Book 1; the rebuilt BVH executes far fewer) / kernel time / peak; see executed_frac",
        "traffic": float(ncu["dram_bytes"]) * share if ncu and share and ncu.get("dram_bytes") else None,
        "kernel": "mega_render_kernel", "kernel_ms": k_ms, "segments_per_launch": seg_local,
        "flop_per_segment": flop_per_seg, "flop_per_segment_source": flop_source,
        "peak_source": "rt_fp32_peak: FP32 FMA-chain microbenchmark on this GPU (MEASURED_PEAKS.json has HBM and "
                       "bf16 tensor peaks only; the path is neither HBM- nor tensor-bound, see DESIGN.md)",
        "hbm": {"achieved": hbm_bytes / (k_ms * 1e-3) * 1e-9, "peak": hbm_peak, "unit": "GB/s",
                "frac": hbm_bytes / (k_ms * 1e-3) * 1e-9 / hbm_peak,
                "achieved_is": "an ESTIMATE from compulsory bytes (scene once + frame once), not a measurement",
                "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "6650 GB/s (of fallback)"},
    }
    if ncu and share:
        # what the kernel EXECUTES (ncu, committed under profiles/): FP32 operations, issue slots, lanes per instruction
        ex_tflops = float(ncu["flop_executed"]) * share / (k_ms * 1e-3) * 1e-12 if ncu.get("flop_executed") else None
        roofline.update({
            "executed": ex_tflops, "executed_frac": ex_tflops / fp32_peak if ex_tflops and fp32_peak > 0 else None,
            "issue_active": ncu.get("issue_active_pct"), "lanes_per_inst": ncu.get("lanes_per_inst"),
            "lane_issue_frac": (ncu["issue_active_pct"] / 100.0 * ncu["lanes_per_inst"] / 32.0
                                if ncu.get("issue_active_pct") and ncu.get("lanes_per_inst") else None),
            "ncu_source": ncu.get("source"), "ncu_scaled_by_segment_share": share,
        })
    else:
        roofline.update({"executed": None, "executed_frac": None, "issue_active": None, "lanes_per_inst": None,
                         "ncu_source": None})
    line = {
        "metric": metric_name(args), "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(args), "kernel": args.kernel, "rng": "per-(pixel,sample) PCG32 streams",
                   "parallelism": f"sample-range x{world}", "l2": "256 MiB buffer zeroed between timed steps",
                   "image": [w, h], "bvh_nodes": info["bvh_nodes"], "scene_bytes": info["device_bytes"],
                   "warmup_spp": warm_spp or args.spp},
        "wall_clock_s": ms_per_step * 1e-3, "segments_per_step": segments,
        "clocks": clocks,
        "reduce_ms": float(red_t.item()), "reduce_bytes": n_values * 4 if world > 1 else 0,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(info["device_bytes"]) + 128,
                "d2h_bytes_per_step": n_values + 48, "ms_per_step": float(e2e_t.item()),
                "ms_steps": [round(v, 2) for v in e2e_ms], "kernel_ms_last": st_e.get("kernel_ms"), "device_total_ms_last": st_e.get("total_ms"),
                "api": "Camera_render (rtc_camera_render_opts): flatten + scene cache lookup (hash of the description; BVH "
                       "build + upload on a miss) + kernels + download"
                       if world == 1 else "rt_scene_acquire + rt_render_device + NCCL reduce + tonemap + D2H per rank"},
        "gpu_launches": launches,
        "roofline": roofline,
        "image_check": image_check,
    }
    if dropin is not None:
        line["e2e_dropin"] = dropin
    if cpu is not None:
        line["cpu_baseline"] = cpu
    print(json.dumps(line))
    renderer.close()
    if world > 1:
        td.destroy_process_group()
    if image_check is not None and image_check["max_abs_diff_u8"] > 1:
        print(f"bench.py: image_check failed: {image_check}", file=sys.stderr)
        sys.exit(3)


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # rank 0 alone runs the CPU reference; the other ranks exit 0 without work
    spp = min(args.ref_spp or CONFIGS[args.config]["ref_spp"], args.spp)
    value, ms, cpu, _ = cpu_reference_arm(args, steps=args.steps, warmup=min(args.warmup, 1), spp=spp)
    line = {
        "impl": "reference", "metric": metric_name(args), "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": min(args.warmup, 1), "ms_per_step": ms, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(args), "step": f"the full frame at {spp} of {args.spp} spp per step "
                                                             f"(cost is linear in spp, raytracing.c:100)"},
        "cpu_baseline": cpu,
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def main():
    args = parse_args()
    if args.impl == "reference":
        reference_arm(args)
    else:
        b200_arm(args)


if __name__ == "__main__":
    main()
