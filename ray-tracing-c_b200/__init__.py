"""B200-native path-tracing backend for the render hot path of gau-nernst/ray-tracing-c.

The product is two in-tree shared libraries (see lib.py); this package is the thin Python mirror used by the
tests, bench.py and the one-process-per-GPU driver.  Import it with
    importlib.import_module("ray-tracing-c_b200")
(the directory name is the project's, and is not a Python identifier).
"""
from . import abi, lib  # noqa: F401
from .render import Scene, Session, fp32_peak_tflops, inspect, make_opts, render_multi  # noqa: F401
