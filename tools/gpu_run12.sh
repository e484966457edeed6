#!/bin/bash
# final 1-GPU measurements of round 2: bench lines of all five configurations, reference arm, launch list, ncu, parity
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv,noheader > gpurun_out/r12_gpu.txt; nproc >> gpurun_out/r12_gpu.txt; grep -m1 "model name" /proc/cpuinfo >> gpurun_out/r12_gpu.txt
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r12_bench_cfg2.json 2> gpurun_out/r12_bench_cfg2.err; echo "cfg2 rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r12_reference_arm.json 2> gpurun_out/r12_reference_arm.err; echo "ref rc=$?"
for c in 1 3 4; do timeout 900 python bench.py --config $c --steps 5 --warmup 3 > gpurun_out/r12_bench_cfg$c.json 2> gpurun_out/r12_bench_cfg$c.err; echo "cfg$c rc=$?"; done
timeout 900 python bench.py --config 5 --steps 1 --warmup 3 --warmup-spp 64 > gpurun_out/r12_bench_cfg5.json 2> gpurun_out/r12_bench_cfg5.err; echo "cfg5 rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r12_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r12_ncu_launch.log 2>&1
for f in book1=1:1200:1000 book2=7:1000:100 cornell=6:600:1000; do
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:mega_render_kernel -c 1 -o gpurun_out/r12_${f%%=*} -f python tools/ab_render.py --frames ${f##*=} --reps 0 > gpurun_out/r12_ncu_${f%%=*}.log 2>&1
done
timeout 1200 python tools/converged_parity.py --configs 1,2,3,4 > gpurun_out/r12_converged.log 2>&1; echo "converged rc=$?"
python - <<'PY'
import json
for c in (1,2,3,4,5):
    try:
        d=json.loads(open(f'gpurun_out/r12_bench_cfg{c}.json').read().strip().splitlines()[-1])
        print(c, round(d['value']), round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value']), round(d['e2e']['ms_per_step'],2), 'frac', round(d['roofline']['frac'],3), 'cpu', round(d['cpu_baseline']['value'],1), d['cpu_baseline']['cores'], 'chk', d['image_check']['max_abs_diff_u8'])
    except Exception as e: print(c, 'ERR', e)
PY
ls -la gpurun_out | grep r12_ | head -40
