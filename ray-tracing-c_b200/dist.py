"""One process per GPU: the image is partitioned by SAMPLE RANGE (BASELINE.json north_star).

Every rank holds the whole (tiny) scene, renders samples [begin, begin+count) of EVERY pixel into a float radiance
buffer with rt_render_device on torch's current CUDA stream, and one torch.distributed reduce (NCCL over
NVLink 5 / NVSwitch, sum, float32, 3*W*H values) brings the buffers to rank 0, which runs the tonemap kernel
(raytracing.c:127-131) and copies the 8-bit image to the host.  The reduce is the path's only exchange step:
9.7 MB for the Book-1 frame, once per render.

PyTorch is plumbing here (device memory, streams, process group); all pixels come from librtc_b200.so.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import abi, lib
from .render import Scene, make_opts


def shard_range(total: int, rank: int, world: int, begin: int = 0) -> tuple[int, int]:
    """Samples [begin + total*rank/world, begin + total*(rank+1)/world): contiguous, disjoint, covering, sizes differ
    by at most one.  Same formula as rt_render_multi (csrc/device/rtc_api.cu)."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    lo = begin + (total * rank) // world
    hi = begin + (total * (rank + 1)) // world
    return lo, hi - lo


def reduce_radiance(radiance, dst: int = 0, group=None):
    """Sum the per-rank radiance buffers onto `dst` (in place).  Works for NCCL (CUDA tensors) and gloo (CPU)."""
    import torch.distributed as dist

    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(radiance, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return radiance


class DistRenderer:
    """Rank-local renderer of one scene.  Usage (under torchrun):

        r = DistRenderer(desc, camera, device=local_rank)
        rgb8, stats = r.render(kernel="mega")       # rgb8 is a numpy array on rank 0, None elsewhere
    """

    def __init__(self, desc: abi.SceneDesc, camera: abi.rt_camera, device: int, group=None, cached: bool = False):
        import torch

        self.torch = torch
        self.device = torch.device("cuda", device)
        torch.cuda.set_device(self.device)
        self.group = group
        self.camera = camera
        self.scene = Scene(desc, device=device, cached=cached)
        h, w = camera.img_height, camera.img_width
        self.radiance = torch.empty((h, w, 3), dtype=torch.float32, device=self.device)
        self.rgb8 = torch.empty((h, w, 3), dtype=torch.uint8, device=self.device)
        self.host_rgb8 = torch.empty((h, w, 3), dtype=torch.uint8, pin_memory=True)
        import torch.distributed as dist

        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1

    def enqueue(self, kernel="mega", rng="sample", drop_nan=False, spp=None):
        """Launch this rank's shard + the reduce + (rank 0) tonemap and D2H on the current stream; asynchronous.
        `spp` renders only the first spp samples of every pixel (warm-up steps of minute-long frames)."""
        spp = int(spp or self.camera.samples_per_pixel)
        begin, count = shard_range(spp, self.rank, self.world)
        if self.world > 1 and rng != "sample":
            raise ValueError("a sample-range shard needs per-sample RNG streams (rng='sample')")
        if count <= 0:
            raise ValueError("fewer samples than ranks")
        opts = make_opts(kernel=kernel, rng=rng, sample_begin=begin, sample_count=count, drop_nan=drop_nan)
        stream = self.torch.cuda.current_stream(self.device).cuda_stream
        self.scene.render_device(self.camera, opts, self.radiance.data_ptr(), stream)
        if self.world > 1:  # the path's one exchange step, event-timed on its own
            self._reduce_ev = (self.torch.cuda.Event(enable_timing=True), self.torch.cuda.Event(enable_timing=True))
            self._reduce_ev[0].record()
        reduce_radiance(self.radiance, 0, self.group)
        if self.world > 1:
            self._reduce_ev[1].record()
        if self.rank == 0:
            rc = lib.b200().rt_tonemap_device(self.radiance.data_ptr(), self.radiance.numel(), spp,
                                              self.rgb8.data_ptr(), stream or None)
            lib.check(rc, "rt_tonemap_device")
            self.host_rgb8.copy_(self.rgb8, non_blocking=True)

    def stats(self) -> dict:
        """Counters of the last enqueue (synchronises the stream).  `reduce_ms`: device time between the end of this
        rank's render kernel and the end of the NCCL reduce on this rank -- the wait for the slowest rank included."""
        stream = self.torch.cuda.current_stream(self.device).cuda_stream
        st = self.scene.device_stats(stream)
        ev = getattr(self, "_reduce_ev", None)
        st["reduce_ms"] = float(ev[0].elapsed_time(ev[1])) if ev is not None else 0.0
        return st

    def render(self, kernel="mega", rng="sample", drop_nan=False, spp=None):
        self.enqueue(kernel, rng, drop_nan, spp)
        st = self.stats()  # synchronises the stream
        image = self.host_rgb8.numpy().copy() if self.rank == 0 else None
        return image, st

    def render_whole(self, kernel="mega", rng="sample"):
        """ALL samples of every pixel on THIS rank's GPU alone (no collective): (rgb8, float radiance sum) as numpy.
        bench.py's image_check compares it with the image the N ranks reduced."""
        torch = self.torch
        spp = int(self.camera.samples_per_pixel)
        whole = torch.empty_like(self.radiance)
        rgb8 = torch.empty_like(self.rgb8)
        stream = torch.cuda.current_stream(self.device).cuda_stream
        opts = make_opts(kernel=kernel, rng=rng, sample_begin=0, sample_count=spp)
        self.scene.render_device(self.camera, opts, whole.data_ptr(), stream)
        lib.check(lib.b200().rt_tonemap_device(whole.data_ptr(), whole.numel(), spp, rgb8.data_ptr(), stream or None),
                  "rt_tonemap_device")
        torch.cuda.synchronize(self.device)
        return rgb8.cpu().numpy(), whole.cpu().numpy()

    def total_segments(self, local_segments: int) -> int:
        """Whole-job segment count (all ranks), on every rank."""
        import torch.distributed as dist

        t = self.torch.tensor([local_segments], dtype=self.torch.int64, device=self.device)
        if dist.is_initialized() and self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return int(t.item())

    def close(self):
        self.scene.close()
