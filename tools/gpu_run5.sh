#!/bin/bash
mkdir -p gpurun_out
python tools/diag_media.py > gpurun_out/r5_diag.log 2>&1; head -12 gpurun_out/r5_diag.log
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/r5_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r5_pytest.log
tail -12 gpurun_out/r5_pytest.log
B1="1:1200:300"; B2="7:1000:100"; C="6:600:300"
ab() { fr=$1; shift; env "$@" timeout 300 python tools/ab_render.py --frames $fr --reps 3 --tag "$*" >> gpurun_out/r5_ab.jsonl 2>> gpurun_out/r5_ab.err; }
ab $B1,$C,$B2 A=default
ab $B2 RTC_BUILD_DIR=$PWD/ray-tracing-c_b200/_build_m0
ab $C RT_MEGA_NV=0 RT_MEGA_MINB=8
ab 0:500:100,2:1200:100,3:1200:100,4:1200:100,5:1200:100 A=default
python - <<'PY'
import json
for l in open('gpurun_out/r5_ab.jsonl'):
    d=json.loads(l); print(f"{d['tag'][-60:]:60s} s{d['scene']} {d['mseg_per_s']:9.1f}")
PY
