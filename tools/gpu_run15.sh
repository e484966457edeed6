#!/bin/bash
mkdir -p gpurun_out
ab() { fr=$1; shift; env "$@" timeout 300 python tools/ab_render.py --frames $fr --reps 3 --tag "$*" >> gpurun_out/r15_ab.jsonl 2>> gpurun_out/r15_ab.err; }
F="1:1200:300,6:600:300,7:1000:100"
ab $F A=default
ab $F RTC_BUILD_DIR=$PWD/ray-tracing-c_b200/_build_v2
ab 1:1200:300 RTC_BUILD_DIR=$PWD/ray-tracing-c_b200/_build_v2 RT_MEGA_THR8=3
ab 1:1200:300 RTC_BUILD_DIR=$PWD/ray-tracing-c_b200/_build_v2 RT_MEGA_THR8=4
ab 1:1200:300 RTC_BUILD_DIR=$PWD/ray-tracing-c_b200/_build_v2 RT_MEGA_THR8=1
python - <<'PY'
import json
for l in open('gpurun_out/r15_ab.jsonl'):
    d=json.loads(l); print(f"{d['tag'][-50:]:50s} s{d['scene']} spp {d['spp']} ms {d['kernel_ms']:.3f} {d['mseg_per_s']:9.1f}")
PY
