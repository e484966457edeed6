#!/bin/bash
mkdir -p gpurun_out
python tools/timeline.py 125 > gpurun_out/r9_timeline.log 2>&1; python tools/timeline.py 1000 >> gpurun_out/r9_timeline.log 2>&1; cat gpurun_out/r9_timeline.log
ab() { fr=$1; shift; env "$@" timeout 300 python tools/ab_render.py --frames $fr --reps 3 --tag "$*" >> gpurun_out/r9_ab.jsonl 2>> gpurun_out/r9_ab.err; }
ab 6:600:300,7:1000:100 A=default
ab 6:600:300,7:1000:100 RT_CONST_CHAINS_OFF=1
ab 6:600:300,7:1000:100 A=default
ab 6:600:300,7:1000:100 RT_CONST_CHAINS_OFF=1
python - <<'PY'
import json
for l in open('gpurun_out/r9_ab.jsonl'):
    d=json.loads(l); print(f"{d['tag'][-40:]:40s} s{d['scene']} spp {d['spp']} ms {d['kernel_ms']:.3f} {d['mseg_per_s']:9.1f}")
PY
