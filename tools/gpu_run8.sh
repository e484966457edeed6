#!/bin/bash
mkdir -p gpurun_out
ab() { fr=$1; shift; env "$@" timeout 300 python tools/ab_render.py --frames $fr --reps 3 --tag "$*" >> gpurun_out/r8_ab.jsonl 2>> gpurun_out/r8_ab.err; }
ab 1:1200:125,1:1200:250,1:1200:1000 A=depth50
abd() { fr=$1; shift; env "$@" timeout 300 python tools/ab_render.py --depth 8 --frames $fr --reps 3 --tag "$*" >> gpurun_out/r8_ab.jsonl 2>> gpurun_out/r8_ab.err; }
abd 1:1200:125,1:1200:250,1:1200:1000 A=depth8
ab 1:1200:125 RT_MEGA_GUIDED=0
ab 1:1200:125 RT_MEGA_CHUNKS=24
ab 1:1200:125 RT_MEGA_CHUNKS=6
python - <<'PY'
import json
for l in open('gpurun_out/r8_ab.jsonl'):
    d=json.loads(l); print(f"{d['tag'][-40:]:40s} s{d['scene']} spp {d['spp']} depth {d['depth']} seg {d['segments']} ms {d['kernel_ms']:.3f} {d['mseg_per_s']:9.1f}")
PY
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mega_render_kernel -c 1 -o gpurun_out/r8_book2 -f python tools/ab_render.py --frames 7:1000:100 --reps 0 > gpurun_out/r8_ncu2.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mega_render_kernel -c 1 -o gpurun_out/r8_cornell -f python tools/ab_render.py --frames 6:600:1000 --reps 0 > gpurun_out/r8_ncu3.log 2>&1
ls -la gpurun_out | grep r8_
