#!/bin/bash
# with the warp item pool: where does a 125-spp launch (one rank of eight) spend its time, and is a quarter still the
# right item fraction?
mkdir -p gpurun_out
for spp in 125 1000; do timeout 120 python tools/timeline.py $spp; done 2>&1 | tee gpurun_out/r30_timeline.log
ab() { fr=$1; shift; env "$@" timeout 100 python tools/ab_render.py --frames $fr --reps 5 --tag "$*" >> gpurun_out/r30_ab.jsonl 2>> gpurun_out/r30_ab.err; }
F="1:1200:125,1:1200:1000,6:600:125"
for g in 4 2 3 6 8; do ab $F RT_MEGA_GUIDED=$g; done
ab $F RT_MEGA_GUIDED=4
python - <<'PY'
import json
rows=[json.loads(l) for l in open('gpurun_out/r30_ab.jsonl')]
for r in rows: print(r['tag'], r['scene'], r['spp'], round(r['kernel_ms'],3), round(r['kernel_ms']/r['spp'],5))
PY
