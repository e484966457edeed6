"""Shared fixtures.  `-m "not gpu"` runs here (no GPU): oracle vs golden vectors, host layer, ABI.  `-m gpu` runs on
a B200: the parity tests proper, through the C ABI."""
from __future__ import annotations

import importlib
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))

GOLDEN = ROOT / "tests" / "golden"
REFERENCE = Path("/root/reference")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def rtc():
    pkg = importlib.import_module("ray-tracing-c_b200")
    build_dir = pkg.lib.BUILD_DIR
    if not (build_dir / "librtc_b200.so").exists() or not (build_dir / "librtc_host.so").exists():
        pkg.lib.build()
    return pkg


@pytest.fixture(scope="session")
def oracle(rtc):
    import oracle_lib

    oracle_lib.oracle()
    return oracle_lib


@pytest.fixture(scope="session")
def ref_tool():
    """The unmodified reference behind oracle/ref_tool.c; only where /root/reference exists (the build container)."""
    tool = ROOT / "oracle" / "_ref" / "ref_tool"
    if not tool.exists():
        if not (REFERENCE / "src" / "main.c").exists():
            pytest.skip("no /root/reference and no prebuilt oracle/_ref/ref_tool")
        subprocess.run(["make", "-C", str(ROOT / "oracle"), "ref"], check=True, capture_output=True)
    return tool


def load_golden(scene_id):
    return np.load(GOLDEN / f"scene{scene_id}.npz")


@pytest.fixture(scope="session")
def golden():
    return load_golden


_session_cache = {}


@pytest.fixture(scope="session")
def scene_factory(rtc):
    """(scene_id, width, spp) -> (Session, SceneDesc); cached, the graphs are never freed anyway."""

    def make(scene_id, width=48, spp=4, **kw):
        key = (scene_id, width, spp, tuple(sorted(kw.items())))
        if key not in _session_cache:
            sess = rtc.Session(scene_id, width=width, spp=spp, **kw)
            _session_cache[key] = (sess, sess.flatten())
        return _session_cache[key]

    return make


def same_bits(a, b):
    """Bitwise equality of float arrays, with any NaN equal to any NaN (x86 and CUDA give different payloads)."""
    a = np.ascontiguousarray(a, np.float32)
    b = np.ascontiguousarray(b, np.float32)
    return (a.view(np.uint32) == b.view(np.uint32)) | (np.isnan(a) & np.isnan(b))


def ulp_distance(a, b):
    a = np.ascontiguousarray(a, np.float32)
    b = np.ascontiguousarray(b, np.float32)
    ia = a.view(np.int32).astype(np.int64)
    ib = b.view(np.int32).astype(np.int64)
    ia = np.where(ia < 0, -(ia & 0x7FFFFFFF), ia)
    ib = np.where(ib < 0, -(ib & 0x7FFFFFFF), ib)
    d = np.abs(ia - ib)
    d[np.isnan(a) & np.isnan(b)] = 0
    return d
