#!/bin/bash
mkdir -p gpurun_out
ab() { fr=$1; shift; env "$@" timeout 300 python tools/ab_render.py --frames $fr --reps 3 --tag "$*" >> gpurun_out/r22_ab.jsonl 2>> gpurun_out/r22_ab.err; }
F="7:1000:100,1:1200:300,6:600:300"
ab $F A=default
ab $F RTC_BUILD_DIR=$PWD/ray-tracing-c_b200/_build_el
ab $F A=default
ab $F RTC_BUILD_DIR=$PWD/ray-tracing-c_b200/_build_el
python - <<'PY'
import json
for l in open('gpurun_out/r22_ab.jsonl'):
    d=json.loads(l); print(f"{d['tag'][-35:]:35s} s{d['scene']} ms {d['kernel_ms']:.3f} {d['mseg_per_s']:9.1f}")
PY
