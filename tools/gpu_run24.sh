#!/bin/bash
# A/B: work items drawn one sample ahead (default build) against -DRTC_ITEM_AHEAD=0 (_build_noahead)
mkdir -p gpurun_out
ab() { fr=$1; shift; env "$@" timeout 300 python tools/ab_render.py --frames $fr --reps 3 --tag "$*" >> gpurun_out/r24_ab.jsonl 2>> gpurun_out/r24_ab.err; }
F="1:1200:32,1:1200:125,1:1200:1000,6:600:125,6:600:1000,7:1000:12,7:1000:100,0:500:100"
ab $F A=ahead
ab $F RTC_BUILD_DIR=/root/repo/ray-tracing-c_b200/_build_noahead
ab $F A=ahead2
python - <<'PY'
import json
rows=[json.loads(l) for l in open('gpurun_out/r24_ab.jsonl')]
for tag in dict.fromkeys(r['tag'] for r in rows):
    print(tag)
    for r in [r for r in rows if r['tag']==tag]:
        print(f"   scene {r['scene']} spp {r['spp']:5d} ms {r['kernel_ms']:9.3f}  ms/spp {r['kernel_ms']/r['spp']:.5f}  Mseg/s {r['mseg_per_s']:9.1f}")
PY
tail -3 gpurun_out/r24_ab.err
