#!/bin/bash
# what the driver runs at round end, on HEAD: the GPU suite with -x, smoke(), the default bench line and its reference arm
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -x -q -m gpu > gpurun_out/final_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/final_pytest.log; tail -4 gpurun_out/final_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 600 python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/final_reference.json 2> gpurun_out/final_reference.err; echo "reference rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/final_bench.json').read().strip().splitlines()[-1])
print(round(d['value']), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value']), 'frac', round(d['roofline']['frac'],3), 'exec_frac', round(d['roofline']['executed_frac'],3), 'lane_issue', round(d['roofline']['lane_issue_frac'],3), 'traffic', d['roofline']['traffic'], 'chk', d['image_check']['max_abs_diff_u8'], 'launches', d['gpu_launches'], 'clocks', d['clocks'])
r=json.loads(open('gpurun_out/final_reference.json').read().strip()); print('reference', round(r['value'],1), r['cpu_baseline']['cores'])
PY
