"""Host-side logic of the multi-GPU path on CPU: sample-range sharding and the reduce, with a world_size-2 gloo
group.  Each rank renders its shard with the ORACLE (test infrastructure; there is no GPU here), the product's
reduce_radiance() sums them, and rank 0 compares with the full-range oracle render."""
from __future__ import annotations

import importlib
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent


def test_shard_range_partitions_the_samples(rtc):
    dist = importlib.import_module("ray-tracing-c_b200.dist")
    for total in (1, 7, 8, 1000, 4096):
        for world in (1, 2, 3, 4, 8):
            if total < world:
                continue
            covered = []
            for rank in range(world):
                lo, n = dist.shard_range(total, rank, world, begin=5)
                assert n >= total // world and n <= total // world + 1
                covered.extend(range(lo, lo + n))
            assert covered == list(range(5, 5 + total))
    with pytest.raises(ValueError):
        dist.shard_range(10, 3, 2)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out_path):
    import torch
    import torch.distributed as td

    sys.path.insert(0, str(ROOT))
    sys.path.insert(0, str(ROOT / "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    td.init_process_group("gloo", rank=rank, world_size=world)
    rtc = importlib.import_module("ray-tracing-c_b200")
    dist = importlib.import_module("ray-tracing-c_b200.dist")
    import oracle_lib as O

    sess = rtc.Session(6, width=24, spp=6)
    desc, cam = sess.flatten(), sess.camera
    lo, n = dist.shard_range(cam.samples_per_pixel, rank, world)
    _, part, _ = O.render(desc, cam, rtc.make_opts(kernel="exact", rng="sample", sample_begin=lo, sample_count=n))
    t = torch.from_numpy(part.copy())
    dist.reduce_radiance(t, dst=0)
    if rank == 0:
        _, whole, _ = O.render(desc, cam, rtc.make_opts(kernel="exact", rng="sample"))
        ok = np.allclose(t.numpy(), whole, rtol=1e-5, atol=1e-6)
        same_image = np.array_equal(O.tonemap(t.numpy(), cam.samples_per_pixel),
                                    O.tonemap(whole, cam.samples_per_pixel))
        Path(out_path).write_text(f"{int(ok)} {int(same_image)}")
    td.barrier()
    td.destroy_process_group()


def test_two_rank_sample_shards_reduce_to_the_full_render(rtc, oracle, tmp_path):
    import torch.multiprocessing as mp

    out = tmp_path / "result.txt"
    mp.spawn(_worker, args=(2, _free_port(), str(out)), nprocs=2, join=True)
    ok, same_image = out.read_text().split()
    assert ok == "1"
    assert same_image == "1"
