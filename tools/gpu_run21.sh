#!/bin/bash
mkdir -p gpurun_out
ab() { fr=$1; shift; env "$@" timeout 300 python tools/ab_render.py --frames $fr --reps 3 --tag "$*" >> gpurun_out/r21_ab.jsonl 2>> gpurun_out/r21_ab.err; }
F="7:1000:100,5:1200:100,3:1200:100"
ab $F A=default
ab $F RTC_BUILD_DIR=$PWD/ray-tracing-c_b200/_build_p2
ab 7:1000:100 RTC_BUILD_DIR=$PWD/ray-tracing-c_b200/_build_p2 RT_MEGA_CTA8=2
ab 7:1000:100 RTC_BUILD_DIR=$PWD/ray-tracing-c_b200/_build_p2 RT_MEGA_THR8=3
python - <<'PY'
import json
for l in open('gpurun_out/r21_ab.jsonl'):
    d=json.loads(l); print(f"{d['tag'][-45:]:45s} s{d['scene']} ms {d['kernel_ms']:.3f} {d['mseg_per_s']:9.1f}")
PY
