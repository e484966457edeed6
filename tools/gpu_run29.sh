#!/bin/bash
# validation + refresh of the measurements after the warp item pool / COLD state (final code of the round)
mkdir -p gpurun_out
timeout 700 python -m pytest tests -x -q -m gpu --timeout 200 > gpurun_out/r29_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r29_pytest.log; tail -4 gpurun_out/r29_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
for c in 2 1 3; do
  timeout 400 python bench.py --config $c --steps 5 --warmup 3 > gpurun_out/r29_bench_cfg$c.json 2> gpurun_out/r29_bench_cfg$c.err; echo "cfg$c rc=$?"
done
timeout 300 python bench.py --config 4 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/r29_bench_cfg4.json 2> gpurun_out/r29_bench_cfg4.err; echo "cfg4 rc=$?"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r29_launches.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-image-check > gpurun_out/r29_ncu_launch.log 2>&1
for f in book1=1:1200:1000 cornell=6:600:1000; do
  timeout 400 ncu --set full --clock-control none --import-source on -k regex:mega_render_kernel -c 1 -o gpurun_out/r29_${f%%=*} -f python tools/ab_render.py --frames ${f##*=} --reps 0 > gpurun_out/r29_ncu_${f%%=*}.log 2>&1; echo "ncu ${f%%=*} rc=$?"
done
python - <<'PY'
import json
for c in (1,2,3,4):
    try:
        d=json.loads(open(f'gpurun_out/r29_bench_cfg{c}.json').read().strip().splitlines()[-1])
        print(c, round(d['value']), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value']), round(d['e2e']['ms_per_step'],2), 'frac', round(d['roofline']['frac'],3), 'cpu', d.get('cpu_baseline') and round(d['cpu_baseline']['value'],1), 'chk', d['image_check'] and d['image_check']['max_abs_diff_u8'], d['clocks'])
    except Exception as e: print(c, 'ERR', e)
PY
