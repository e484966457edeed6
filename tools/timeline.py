#!/usr/bin/env python
"""Development: with a -DRTC_TIMELINE build (RTC_BUILD_DIR) every warp of mega_render_kernel prints when it started,
when it first found the item queue empty, when it last drew items (and of which chunk), and when it ended.  Summarises
where a launch's time goes and what the last warps were doing."""
import os, subprocess, sys, numpy as np
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
spp = sys.argv[1] if len(sys.argv) > 1 else "125"
frame = sys.argv[2] if len(sys.argv) > 2 else "1:1200"
env = dict(os.environ, RTC_BUILD_DIR=str(ROOT / "ray-tracing-c_b200" / "_build_tl"))
out = subprocess.run([sys.executable, str(ROOT / "tools" / "ab_render.py"), "--frames", f"{frame}:{spp}", "--reps", "1"],
                     env=env, capture_output=True, text=True).stdout
rows = [l.split()[1:] for l in out.splitlines() if l.startswith("TL ")]
a = np.array(rows, dtype=np.float64)
n = len(a) // 2  # two renders (warm-up + timed): keep the second
a = a[n:]
t0 = a[:, 1].min()
ms = lambda col: (a[:, col] - t0) * 1e-6
start, dry, end, trips, last, chunk, draws, sm = ms(1), ms(2), ms(3), a[:, 4], ms(5), a[:, 6], a[:, 7], a[:, 8]
q = lambda v, ps=(0, 50, 90, 99, 100): " ".join(f"{np.percentile(v, p):.3f}" for p in ps)
print(f"spp {spp}: {len(a)} warps; start spread {start.max():.3f} ms")
print(f"  queue first seen empty (ms)   min/p50/p90/p99/max: {q(dry[a[:, 2] > 0])}")
print(f"  warp's last draw (ms)         min/p50/p90/p99/max: {q(last)}   chunk of it: min {chunk.min():.0f} median {np.median(chunk):.0f} max {chunk.max():.0f}")
print(f"  warp end (ms)                 min/p50/p90/p99/max: {q(end)}")
print(f"  end - last draw (ms)          min/p50/p90/p99/max: {q(end - last)}")
print(f"  draws per warp                min/p50/p90/p99/max: {q(draws)};  trips {q(trips)}")
late = np.argsort(end)[-12:]
print("  the 12 last warps: end, last draw, chunk of it, draws, trips, sm")
for i in late:
    print(f"    {end[i]:8.3f} {last[i]:8.3f} {chunk[i]:4.0f} {draws[i]:6.0f} {trips[i]:6.0f} {sm[i]:4.0f}")
per_sm = {int(s): end[sm == s].max() for s in np.unique(sm)}
v = np.array(list(per_sm.values()))
print(f"  last warp end per SM (ms)     min/p50/p90/p99/max: {q(v)}")
hist, edges = np.histogram(end, bins=12)
print("  warp-end histogram:", " ".join(f"{edges[i]:.2f}:{hist[i]}" for i in range(len(hist))))
print([l for l in out.splitlines() if l.startswith("{")][-1][:300])
