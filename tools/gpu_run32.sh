#!/bin/bash
# A/B: pool batch decoded once per draw (default build) against once per item (_build_prev)
mkdir -p gpurun_out
echo prev; RTC_BUILD_DIR=/root/repo/ray-tracing-c_b200/_build_prev timeout 60 python tools/dev/pool_check.py 2>/dev/null
echo new; timeout 60 python tools/dev/pool_check.py 2>/dev/null; echo rc=$?
ab() { fr=$1; shift; env "$@" timeout 100 python tools/ab_render.py --frames $fr --reps 5 --tag "$*" >> gpurun_out/r32_ab.jsonl 2>> gpurun_out/r32_ab.err; }
F="0:500:100,1:1200:32,1:1200:125,1:1200:1000,6:600:125,6:600:1000"
for rep in 1 2; do
ab $F A=default
ab $F RTC_BUILD_DIR=/root/repo/ray-tracing-c_b200/_build_prev
done
python - <<'PY'
import json
rows=[json.loads(l) for l in open('gpurun_out/r32_ab.jsonl')]
tags=list(dict.fromkeys(r['tag'] for r in rows))
keys=list(dict.fromkeys((r['scene'],r['spp']) for r in rows))
print('scene spp  '+'  '.join(t[-12:] for t in tags))
for k in keys:
    v=[min([r['kernel_ms'] for r in rows if r['tag']==t and (r['scene'],r['spp'])==k] or [float('nan')]) for t in tags]
    sh=[[r['sha'] for r in rows if r['tag']==t and (r['scene'],r['spp'])==k][:1] for t in tags]
    print(k, '  '.join(f"{x:9.3f}" for x in v), '  '.join(f"{100*(x/v[0]-1):+.2f}%" for x in v[1:]), 'same image' if sh[0]==sh[1] else sh)
PY
