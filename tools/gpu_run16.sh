#!/bin/bash
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r16_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r16_pytest.log
tail -6 gpurun_out/r16_pytest.log
ab() { fr=$1; shift; env "$@" timeout 300 python tools/ab_render.py --frames $fr --reps 3 --tag "$*" >> gpurun_out/r16_ab.jsonl 2>> gpurun_out/r16_ab.err; }
F="1:1200:300,6:600:300,7:1000:100,0:500:100,2:1200:100,4:1200:100,5:1200:100"
ab $F A=default
ab $F RT_BVH_PRE=0
ab 1:1200:300 RT_MEGA_NV=-1
python - <<'PY'
import json
for l in open('gpurun_out/r16_ab.jsonl'):
    d=json.loads(l); print(f"{d['tag'][-50:]:50s} s{d['scene']} spp {d['spp']} ms {d['kernel_ms']:.3f} {d['mseg_per_s']:9.1f}")
PY
