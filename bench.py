#!/usr/bin/env python
"""bench.py -- throughput of the per-pixel render loop on B200s, beside the reference's OpenMP build.

    python bench.py --gpus N --steps K --warmup W            this backend (librtc_b200.so through the C ABI)
    python bench.py --impl reference --gpus N --steps K ...  the UNMODIFIED reference (oracle/_ref/ref_main_fast,
                                                             -Ofast -fopenmp, all host cores) on the same config

Workload (BASELINE.json configs[1], the configuration the metric is quoted on): the Book 1 final scene
(`./main 1 1200 1000 x`): 486 spheres, 1200x675, 1000 samples per pixel, depth 50.  One STEP = one whole render of
that frame.  For N > 1 (launched by torchrun, one rank per GPU) the 1000 samples of every pixel are split into N
contiguous ranges (strong scaling: the frame is fixed), each rank renders its range into a float radiance buffer,
one NCCL sum-reduce brings them to rank 0, which tonemaps.

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

METRIC = "path segments per second (Mrays/s), Book 1 final scene 1200x675, 1000 spp, depth 50"
UNIT = "Mseg/s"
SCENE_ID, WIDTH, SPP, DEPTH = 1, 1200, 1000, 50
REF_FAST = ROOT / "oracle" / "_ref" / "ref_main_fast"
# dram__bytes_read.sum + dram__bytes_write.sum of mega_render_kernel per launch, from profiles/ (ncu --set full);
# profiles/r01j_mega_render_kernel_book1_1000spp.txt (the bench's own launch: 1200x675, 1000 spp): 27.5 MB read +
# 5.062 GB written -- local-memory (traversal stack, spills) write-backs from L2, 22 GB/s, 0.26 % of DRAM peak
NCU_DRAM_BYTES_PER_LAUNCH = 5.090e9
FALLBACK_FLOP_PER_SEGMENT = 1434.0  # scene 1; recomputed live from the oracle's counters when possible


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--kernel", default="mega", choices=["mega", "exact", "wave"])
    # smaller frames for development only; a line produced with these says so in config.workload
    ap.add_argument("--scene", type=int, default=SCENE_ID)
    ap.add_argument("--width", type=int, default=WIDTH)
    ap.add_argument("--spp", type=int, default=SPP)
    ap.add_argument("--ref-spp", type=int, default=0, help="samples per pixel of one reference/CPU step (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def workload_name(args):
    titles = {0: "scene 0 (metal+lambertian)", 1: "Book 1 final scene (486 spheres)", 6: "Cornell box (mixture pdf)",
              7: "Book 2 final scene"}
    name = titles.get(args.scene, f"scene {args.scene}")
    return f"{name}, width {args.width}, {args.spp} spp, depth {DEPTH}"


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
    sess = rtc.Session(args.scene, width=args.width, spp=2, depth=DEPTH)
    desc = sess.flatten()
    _, _, st = O.render(desc, sess.camera, rtc.make_opts(kernel="exact", rng="pixel"), O.MATH_LIBM)
    work = O.last_work()
    return st["segments"] / st["samples"], O.flop_per_segment(work), work


def run_reference_once(args, spp: int, threads: int) -> float:
    """Wall seconds of `./main <scene> <width> <spp> x` of the unmodified reference (scene build < 0.1 s and the TIFF
    write are inside, as in the reference's own timing of the process)."""
    env = dict(os.environ, OMP_NUM_THREADS=str(threads))
    with tempfile.TemporaryDirectory() as tmp:
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
        return time.perf_counter() - t0


def run_port_once(args, spp: int) -> float:
    """Fallback when the reference binary is absent: the oracle port (OpenMP over rows)."""
    sys.path.insert(0, str(ROOT / "tests"))
    import oracle_lib as O

    rtc = O.rtc
    sess = rtc.Session(args.scene, width=args.width, spp=spp, depth=DEPTH)
    desc = sess.flatten()
    t0 = time.perf_counter()
    O.render(desc, sess.camera, rtc.make_opts(kernel="exact", rng="pixel"), O.MATH_LIBM)
    return time.perf_counter() - t0


def pixels_of(args) -> int:
    rtc = importlib.import_module("ray-tracing-c_b200")
    sess = rtc.Session(args.scene, width=args.width, spp=1, depth=DEPTH)
    return sess.width * sess.height


def cpu_reference_arm(args, steps: int, warmup: int, spp: int):
    """Times `steps` bounded samples of the workload on the host cores.  Returns (Mseg/s, ms per step, descriptor)."""
    cores = host_cores()
    seg_per_sample, flop_per_seg, _ = oracle_segments_per_sample(args)
    kind = "reference" if REF_FAST.exists() else "port"
    run = (lambda: run_reference_once(args, spp, cores)) if kind == "reference" else (lambda: run_port_once(args, spp))
    for _ in range(warmup):
        run()
    times = [run() for _ in range(max(steps, 1))]
    pixels = pixels_of(args)
    segments = pixels * spp * seg_per_sample
    sec = float(np.mean(times))
    value = segments / sec * 1e-6
    desc = {
        "value": value, "unit": UNIT, "cores": cores, "kind": kind,
        "sample": (f"{'oracle/_ref/ref_main_fast (unmodified reference, gcc -Ofast -fopenmp)' if kind == 'reference' else 'oracle port (liboracle.so, OpenMP)'}"
                   f" `main {args.scene} {args.width} {spp} x`: the full {args.width}-wide frame at {spp} of {args.spp} spp,"
                   f" {sec:.2f} s per run, mean of {len(times)}; segments = pixels*spp*{seg_per_sample:.3f} seg/sample"
                   f" counted by the oracle on the same frame"),
        "seconds_per_run": sec, "spp": spp, "seg_per_sample": seg_per_sample,
        "full_frame_seconds_extrapolated": sec * args.spp / spp,
    }
    return value, sec * 1e3, desc, flop_per_seg


# ---------------------------------------------------------------------- this backend --------------------------------

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
    if world > 1:
        td.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    rtc = importlib.import_module("ray-tracing-c_b200")
    dist = importlib.import_module("ray-tracing-c_b200.dist")
    sess = rtc.Session(args.scene, width=args.width, spp=args.spp, depth=DEPTH)
    desc, cam = sess.flatten(), sess.camera
    info = rtc.inspect(desc)
    renderer = dist.DistRenderer(desc, cam, device=local_rank)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")  # > 126 MB L2

    def barrier():
        if world > 1:
            td.barrier()
        torch.cuda.synchronize()

    # ---- resident-scene throughput (`value`)
    for _ in range(args.warmup):
        renderer.enqueue(kernel=args.kernel)
        torch.cuda.synchronize()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    barrier()
    step_ms, kernel_ms, seg_local, launches = [], [], 0, 0
    for _ in range(args.steps):
        flush.zero_()  # L2 flush between timed iterations (untimed)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        renderer.enqueue(kernel=args.kernel)
        e1.record()
        st = renderer.stats()  # synchronises; kernel_ms comes from CUDA events around the kernel on its own stream
        step_ms.append(e0.elapsed_time(e1))
        kernel_ms.append(st["kernel_ms"])
        seg_local = st["segments"]
        launches += st["launches"] + (1 if rank == 0 else 0)  # + the tonemap on rank 0
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    total_ms = torch.tensor([sum(step_ms)], dtype=torch.float64, device="cuda")
    if world > 1:
        td.all_reduce(total_ms, op=td.ReduceOp.MAX)
    ms_per_step = float(total_ms.item()) / args.steps
    segments = renderer.total_segments(seg_local)
    value = segments / ms_per_step * 1e-3

    # ---- end to end through the reference-facing call, host buffers in and out (`e2e`)
    e2e_steps = min(args.steps, 3)
    h, w = cam.img_height, cam.img_width
    barrier()
    e2e_ms = []
    e2e_warm = max(1, min(args.warmup, 3))
    for it in range(-e2e_warm, e2e_steps):  # negative iterations: untimed warm-up (first-use allocations of the pools)
        barrier()
        t0 = time.perf_counter()
        if world == 1:
            # Camera_render itself: World/Camera structs in host memory -> uint8 image in host memory
            rgb8, _, st_e = sess.camera_render(rtc.make_opts(kernel=args.kernel, rng="sample"))
        else:
            r2 = dist.DistRenderer(desc, cam, device=local_rank)  # uploads the scene from the host description
            rgb8, st_e = r2.render(kernel=args.kernel)
            r2.close()
        barrier()
        if it >= 0:
            e2e_ms.append((time.perf_counter() - t0) * 1e3)
    e2e_t = torch.tensor([float(np.mean(e2e_ms))], dtype=torch.float64, device="cuda")
    if world > 1:
        td.all_reduce(e2e_t, op=td.ReduceOp.MAX)
    e2e_value = segments / float(e2e_t.item()) * 1e-3

    if rank != 0:
        renderer.close()
        if world > 1:
            td.destroy_process_group()
        return

    # ---- roofline of the dominant kernel (mega_render_kernel): FP32 issue, not HBM and not tensor
    fp32_peak = rtc.fp32_peak_tflops(local_rank)
    cpu = None
    flop_per_seg = FALLBACK_FLOP_PER_SEGMENT
    if world == 1 and not args.no_cpu_baseline:
        ref_spp = args.ref_spp or 32
        _, _, cpu, flop_per_seg = cpu_reference_arm(args, steps=1, warmup=0, spp=ref_spp)
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
    # sums, 12 frames for this configuration, live in L2-sized buffers and are an implementation choice)
    hbm_bytes = n_values * 4 + int(info["device_bytes"])
    roofline = {
        "bound": "fp32", "achieved": achieved_tflops, "peak": fp32_peak, "unit": "TFLOP/s",
        "frac": achieved_tflops / fp32_peak if fp32_peak > 0 else None,
        "traffic": NCU_DRAM_BYTES_PER_LAUNCH,
        "kernel": "mega_render_kernel", "kernel_ms": k_ms, "segments_per_launch": seg_local,
        "flop_per_segment": flop_per_seg,
        "peak_source": "rt_fp32_peak: FP32 FMA-chain microbenchmark on this GPU (MEASURED_PEAKS.json has HBM and "
                       "bf16 tensor peaks only; the path is neither HBM- nor tensor-bound, see DESIGN.md)",
        "hbm": {"achieved": hbm_bytes / (k_ms * 1e-3) * 1e-9, "peak": hbm_peak, "unit": "GB/s",
                "frac": hbm_bytes / (k_ms * 1e-3) * 1e-9 / hbm_peak,
                "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "6650 GB/s (of fallback)"},
    }
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(args), "kernel": args.kernel, "rng": "per-(pixel,sample) PCG32 streams",
                   "parallelism": f"sample-range x{world}", "l2": "256 MiB buffer zeroed between timed steps",
                   "image": [w, h], "bvh_nodes": info["bvh_nodes"], "scene_bytes": info["device_bytes"]},
        "wall_clock_s": ms_per_step * 1e-3, "segments_per_step": segments,
        "clocks": clocks,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(info["device_bytes"]) + 128,
                "d2h_bytes_per_step": n_values + 48, "ms_per_step": float(e2e_t.item()),
                "ms_steps": [round(v, 2) for v in e2e_ms], "kernel_ms_last": st_e.get("kernel_ms"), "device_total_ms_last": st_e.get("total_ms"),
                "api": "Camera_render (rtc_camera_render_opts): flatten + BVH build + upload + kernels + download"
                       if world == 1 else "rt_scene_create + rt_render_device + NCCL reduce + tonemap + D2H per rank"},
        "gpu_launches": launches,
        "roofline": roofline,
    }
    if cpu is not None:
        line["cpu_baseline"] = cpu
    print(json.dumps(line))
    renderer.close()
    if world > 1:
        td.destroy_process_group()


def reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return  # rank 0 alone runs the CPU reference; the other ranks exit 0 without work
    spp = args.ref_spp or 16
    value, ms, cpu, _ = cpu_reference_arm(args, steps=args.steps, warmup=min(args.warmup, 1), spp=spp)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
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
