#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_mega.py -m gpu -q -k "build_options or variants" > gpurun_out/r13_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r13_pytest.log
tail -15 gpurun_out/r13_pytest.log
ab() { fr=$1; shift; env "$@" timeout 300 python tools/ab_render.py --frames $fr --reps 3 --tag "$*" >> gpurun_out/r13_ab.jsonl 2>> gpurun_out/r13_ab.err; }
F="1:1200:300,6:600:300,7:1000:100,0:500:100,5:1200:100"
ab $F A=default
ab $F RT_BVH_WIDE=1
ab 7:1000:100 RT_BVH_WIDE=1 RT_MEGA_NV=0 RT_MEGA_MINB=8
ab 1:1200:300 RT_BVH_WIDE=1 RT_MEGA_THR8=1
ab 1:1200:300 RT_BVH_WIDE=1 RT_MEGA_THR8=3
ab 1:1200:300 RT_BVH_WIDE=1 RT_MEGA_THR8=4
python - <<'PY'
import json
for l in open('gpurun_out/r13_ab.jsonl'):
    d=json.loads(l); print(f"{d['tag'][-50:]:50s} s{d['scene']} spp {d['spp']} ms {d['kernel_ms']:.3f} {d['mseg_per_s']:9.1f}")
PY
python -c "
import json; d=json.loads(open('gpurun_out/.x','w') and '{}')" 2>/dev/null
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
