#!/bin/bash
mkdir -p gpurun_out
ab() { fr=$1; shift; env "$@" timeout 300 python tools/ab_render.py --frames $fr --reps 3 --tag "$*" >> gpurun_out/r23_ab.jsonl 2>> gpurun_out/r23_ab.err; }
F="1:1200:32,1:1200:62,1:1200:125,1:1200:250,1:1200:500,1:1200:1000"
ab $F A=default
ab $F RT_MEGA_GUIDED=0
env timeout 300 python tools/ab_render.py --depth 4 --frames $F --reps 3 --tag depth4 >> gpurun_out/r23_ab.jsonl
python - <<'PY'
import json
rows=[json.loads(l) for l in open('gpurun_out/r23_ab.jsonl')]
for tag in sorted(set(r['tag'] for r in rows)):
    rs=[r for r in rows if r['tag']==tag]
    print(tag)
    for r in rs: print(f"   spp {r['spp']:5d} ms {r['kernel_ms']:9.3f}  ms/spp {r['kernel_ms']/r['spp']:.5f}  Mseg/s {r['mseg_per_s']:9.1f}")
    a,b=rs[-2],rs[-1]; slope=(b['kernel_ms']-a['kernel_ms'])/(b['spp']-a['spp'])
    print('   slope', round(slope,5), 'fixed:', [round(r['kernel_ms']-slope*r['spp'],3) for r in rs])
PY
