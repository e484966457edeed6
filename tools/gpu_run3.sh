#!/bin/bash
# round 2, GPU call 3: full GPU test suite, A/B of variants and knobs
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r3_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r3_pytest.log
tail -15 gpurun_out/r3_pytest.log
B1="1:1200:300"; B2="7:1000:100"; C="6:600:300"
ab() { fr=$1; shift; env "$@" timeout 300 python tools/ab_render.py --frames $fr --reps 3 --tag "$*" >> gpurun_out/r3_ab.jsonl 2>> gpurun_out/r3_ab.err; }
ab $B1,$C,$B2 A=default
ab $C,$B2 RT_BVH_BOXES=1
ab $B1,$C,$B2 RTC_BUILD_DIR=$PWD/ray-tracing-c_b200/_build_d0
ab $B2 RTC_BUILD_DIR=$PWD/ray-tracing-c_b200/_build_m0
ab $B1 RT_MEGA_MINB=10
ab $B1 RT_MEGA_MINB=12
ab $B1 RT_MEGA_THR8=1
ab $B1 RT_MEGA_THR8=3
ab $B1 RT_MEGA_CHUNKS=24
ab $B2 RT_MEGA_NV=-2 RT_MEGA_MINB=4
ab $B2 RT_MEGA_NV=-1 RT_MEGA_MINB=2
ab $B2 RT_MEGA_NV=-1 RT_MEGA_MINB=4
ab $B2 RT_MEGA_NV=-3 RT_MEGA_MINB=8
ab $B2 RT_MEGA_NV=0 RT_MEGA_MINB=8
ab $B2 RT_MEGA_CTA8=4
ab $B2 RT_MEGA_THR8=3
ab $C RT_MEGA_NV=0 RT_MEGA_MINB=8
ab $C RT_MEGA_NV=-1 RT_MEGA_MINB=2
python - <<'PY'
import json
for l in open('gpurun_out/r3_ab.jsonl'):
    d=json.loads(l); print(f"{d['tag'][-60:]:60s} s{d['scene']} {d['mseg_per_s']:9.1f}")
PY
