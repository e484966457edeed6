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


def test_bench_reference_arm_contract(tmp_path):
    """`bench.py --impl reference` (the driver's reference arm): rank 0 prints ONE JSON line with the metric of the
    backend's own arm, `impl: reference`, a cpu_baseline describing the run and an e2e object with zero copies;
    every other rank exits 0 without work.  Small frame: the arm times the reference's CPU code, no GPU involved."""
    import json
    import os
    import subprocess
    import sys
    from pathlib import Path

    root = Path(__file__).resolve().parent.parent
    cmd = [sys.executable, str(root / "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0",
           "--width", "160", "--spp", "8", "--ref-spp", "2"]
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "LOCAL_RANK", "WORLD_SIZE")}
    out = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    line = json.loads(lines[0])
    assert line["impl"] == "reference" and line["unit"] == "Mseg/s" and line["higher_is_better"] is True
    assert line["value"] > 0 and line["ms_per_step"] > 0 and line["vs_baseline"] is None
    assert line["cpu_baseline"]["kind"] in ("reference", "port") and line["cpu_baseline"]["cores"] >= 1
    assert line["cpu_baseline"]["value"] == line["value"] and line["cpu_baseline"]["sample"]
    assert line["e2e"] == {"value": line["value"], "unit": "Mseg/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    other = subprocess.run(cmd, env=dict(env, RANK="1", LOCAL_RANK="1", WORLD_SIZE="2"), capture_output=True,
                           text=True, timeout=300)
    assert other.returncode == 0 and other.stdout.strip() == ""
