#!/bin/bash
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/r10_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r10_pytest.log
tail -6 gpurun_out/r10_pytest.log
python tools/timeline.py 125 > gpurun_out/r10_timeline.log 2>&1; cat gpurun_out/r10_timeline.log | head -3
ab() { fr=$1; shift; env "$@" timeout 300 python tools/ab_render.py --frames $fr --reps 3 --tag "$*" >> gpurun_out/r10_ab.jsonl 2>> gpurun_out/r10_ab.err; }
ab 1:1200:125,1:1200:1000,6:600:125,6:600:1000,7:1000:100,0:500:100 A=default
ab 1:1200:125,1:1200:1000,6:600:125,7:1000:100 RT_MEGA_GUIDED=0
ab 7:1000:100 RT_MEGA_MINB=3
python - <<'PY'
import json
for l in open('gpurun_out/r10_ab.jsonl'):
    d=json.loads(l); print(f"{d['tag'][-40:]:40s} s{d['scene']} spp {d['spp']} ms {d['kernel_ms']:.3f} {d['mseg_per_s']:9.1f}")
PY
