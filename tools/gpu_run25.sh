#!/bin/bash
# A/B/C: items drawn one sample ahead always (default build), never (_build_noahead), only from items of >= 4 samples
mkdir -p gpurun_out
ab() { fr=$1; shift; env "$@" timeout 300 python tools/ab_render.py --frames $fr --reps 5 --tag "$*" >> gpurun_out/r25_ab.jsonl 2>> gpurun_out/r25_ab.err; }
F="0:500:100,0:500:1000,1:1200:125,1:1200:1000,6:600:125,6:600:1000,7:1000:25,7:1000:100,5:800:200,4:400:200,2:400:200"
for rep in 1 2; do
ab $F A=ahead
ab $F RTC_BUILD_DIR=/root/repo/ray-tracing-c_b200/_build_noahead
ab $F RTC_BUILD_DIR=/root/repo/ray-tracing-c_b200/_build_ahead4
done
python - <<'PY'
import json
rows=[json.loads(l) for l in open('gpurun_out/r25_ab.jsonl')]
tags=list(dict.fromkeys(r['tag'] for r in rows))
keys=list(dict.fromkeys((r['scene'],r['spp']) for r in rows))
print('scene spp  '+'  '.join(t[-12:] for t in tags))
for k in keys:
    print(k, '  '.join(f"{min(r['kernel_ms'] for r in rows if r['tag']==t and (r['scene'],r['spp'])==k):9.3f}" for t in tags))
PY
tail -3 gpurun_out/r25_ab.err
