#!/bin/bash
# A/B: per-lane state that no traversal touches (sample sums, counters) in shared memory (-DRTC_COLD_SMEM=1)
mkdir -p gpurun_out
ab() { fr=$1; shift; env "$@" timeout 300 python tools/ab_render.py --frames $fr --reps 5 --tag "$*" >> gpurun_out/r26_ab.jsonl 2>> gpurun_out/r26_ab.err; }
F="0:500:100,0:500:1000,1:1200:125,1:1200:1000,6:600:125,6:600:1000,7:1000:25,7:1000:100,5:800:200,4:400:200,2:400:200,3:400:200"
for rep in 1 2; do
ab $F A=default
ab $F RTC_BUILD_DIR=/root/repo/ray-tracing-c_b200/_build_cold
done
python - <<'PY'
import json
rows=[json.loads(l) for l in open('gpurun_out/r26_ab.jsonl')]
tags=list(dict.fromkeys(r['tag'] for r in rows))
keys=list(dict.fromkeys((r['scene'],r['spp']) for r in rows))
print('scene spp  '+'  '.join(t[-12:] for t in tags))
for k in keys:
    v=[min(r['kernel_ms'] for r in rows if r['tag']==t and (r['scene'],r['spp'])==k) for t in tags]
    print(k, '  '.join(f"{x:9.3f}" for x in v), f"{100*(v[1]/v[0]-1):+.2f}%")
PY
tail -3 gpurun_out/r26_ab.err
