#!/bin/bash
mkdir -p gpurun_out
ab() { fr=$1; shift; env "$@" timeout 300 python tools/ab_render.py --frames $fr --reps 3 --tag "$*" >> gpurun_out/r17_ab.jsonl 2>> gpurun_out/r17_ab.err; }
B1="1:1200:300"; C="6:600:300"; B2="7:1000:100"
ab $B1,$C,$B2 A=default
for t in 1 3 4; do ab $B1,$C,$B2 RT_MEGA_THR8=$t; done
for c in 2 4 5 6; do ab $B2 RT_MEGA_CTA8=$c; done
ab $B2 RT_MEGA_RESUME8=2
ab $B2 RT_MEGA_RESUME8=0
ab $C RT_MEGA_MINB=6
ab $C RT_MEGA_MINB=4
ab $B1 RT_MEGA_MINB=6
python - <<'PY'
import json
for l in open('gpurun_out/r17_ab.jsonl'):
    d=json.loads(l); print(f"{d['tag'][-40:]:40s} s{d['scene']} ms {d['kernel_ms']:.3f} {d['mseg_per_s']:9.1f}")
PY
