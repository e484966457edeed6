#!/bin/bash
# refresh of the Book 1 / scene 0 measurements after the FEAT_PRE kernels (final code)
mkdir -p gpurun_out
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r18_bench_cfg2.json 2> gpurun_out/r18_bench_cfg2.err; echo "cfg2 rc=$?"
timeout 600 python bench.py --config 1 --steps 5 --warmup 3 > gpurun_out/r18_bench_cfg1.json 2> gpurun_out/r18_bench_cfg1.err; echo "cfg1 rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r18_reference_arm.json 2> gpurun_out/r18_reference_arm.err; echo "ref rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r18_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r18_ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mega_render_kernel -c 1 -o gpurun_out/r18_book1 -f python tools/ab_render.py --frames 1:1200:1000 --reps 0 > gpurun_out/r18_ncu_book1.log 2>&1
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
python - <<'PY'
import json
for c in (1,2):
    d=json.loads(open(f'gpurun_out/r18_bench_cfg{c}.json').read().strip().splitlines()[-1])
    print(c, round(d['value']), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value']), round(d['e2e']['ms_per_step'],2), 'frac', round(d['roofline']['frac'],3), 'exec', d['roofline'].get('executed_frac'), 'cpu', round(d['cpu_baseline']['value'],1), d['cpu_baseline']['cores'], 'chk', d['image_check']['max_abs_diff_u8'])
PY
