#!/usr/bin/env python
"""Development: with a -DRTC_TIMELINE build (RTC_BUILD_DIR) every CTA of mega_render_kernel prints when it started, when
its first lane found the item queue empty, and when it ended.  Summarises where a launch's time goes."""
import os, subprocess, sys, numpy as np
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
spp = sys.argv[1] if len(sys.argv) > 1 else "125"
env = dict(os.environ, RTC_BUILD_DIR=str(ROOT / "ray-tracing-c_b200" / "_build_tl"))
out = subprocess.run([sys.executable, str(ROOT / "tools" / "ab_render.py"), "--frames", f"1:1200:{spp}", "--reps", "1"],
                     env=env, capture_output=True, text=True).stdout
rows = [l.split()[1:] for l in out.splitlines() if l.startswith("TL ")]
a = np.array(rows, dtype=np.float64)
n = len(a) // 2  # two renders (warm-up + timed): keep the second
a = a[n:]
t0 = a[:, 1].min()
start, dry, end, trips = (a[:, 1] - t0) * 1e-6, (a[:, 2] - t0) * 1e-6, (a[:, 3] - t0) * 1e-6, a[:, 4]
print(f"spp {spp}: {len(a)} CTAs; start spread {start.max():.3f} ms; first lane dry: min {dry.min():.3f} median {np.median(dry):.3f} max {dry.max():.3f} ms;"
      f" CTA end: min {end.min():.3f} median {np.median(end):.3f} p99 {np.percentile(end, 99):.3f} max {end.max():.3f} ms; trips median {np.median(trips):.0f} max {trips.max():.0f}")
print([l for l in out.splitlines() if l.startswith("{")][-1][:300])
