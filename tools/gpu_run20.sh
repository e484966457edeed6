#!/bin/bash
mkdir -p gpurun_out
ab() { fr=$1; shift; env "$@" timeout 300 python tools/ab_render.py --frames $fr --reps 3 --tag "$*" >> gpurun_out/r20_ab.jsonl 2>> gpurun_out/r20_ab.err; }
F="1:1200:125,1:1200:1000,6:600:125,7:1000:125"
for g in 4 3 6 8 12; do ab $F RT_MEGA_GUIDED=$g; done
python - <<'PY'
import json
for l in open('gpurun_out/r20_ab.jsonl'):
    d=json.loads(l); print(f"{d['tag'][-30:]:30s} s{d['scene']} spp {d['spp']} ms {d['kernel_ms']:.3f} {d['mseg_per_s']:9.1f}")
PY
