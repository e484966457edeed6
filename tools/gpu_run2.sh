#!/bin/bash
# round 2, GPU call 2: tests (F10 exact media, box leaves, lean kernels), A/B of this round's switches, bench, parity
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest.log
tail -15 gpurun_out/r2_pytest.log
F="1:1200:300,6:600:300,7:1000:100"
ab() { env "$@" timeout 300 python tools/ab_render.py --frames $F --reps 3 --tag "$*" >> gpurun_out/r2_ab.jsonl 2>> gpurun_out/r2_ab.err; }
ab A=default
ab RT_MEGA_LEAN=0
ab RT_BVH_BOXES=0
ab RT_ONE_LIGHT_OFF=1
ab RT_MEGA_GUIDED=0
ab RT_BVH_BOXES=0 RT_ONE_LIGHT_OFF=1 RT_MEGA_LEAN=0 RT_MEGA_GUIDED=0
ab RT_BVH_COST_BOX=2.5
ab RT_BVH_COST_BOX=6
cat gpurun_out/r2_ab.jsonl | cut -c1-400
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err; echo "bench rc=$?"
timeout 900 python tools/converged_parity.py --configs 2,3,4 > gpurun_out/r2_converged.log 2>&1; echo "converged rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mega_render_kernel -c 1 -o gpurun_out/r2_book1 -f python tools/ab_render.py --frames 1:1200:1000 --reps 0 > gpurun_out/r2_ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mega_render_kernel -c 1 -o gpurun_out/r2_book2 -f python tools/ab_render.py --frames 7:1000:100 --reps 0 > gpurun_out/r2_ncu2.log 2>&1
ls -la gpurun_out/ | tail -12
