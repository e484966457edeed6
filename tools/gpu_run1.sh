#!/bin/bash
# round 2, GPU call 1: tests, bench, shared-memory-stack A/B, ncu of the bench launch
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv > gpurun_out/r1_gpu.txt; nproc >> gpurun_out/r1_gpu.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r1_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r1_pytest.log
tail -5 gpurun_out/r1_pytest.log
timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/r1_bench.json 2> gpurun_out/r1_bench.err; echo "bench rc=$?"
for b in _build_s0 _build _build_s12 _build_s16; do
  RTC_BUILD_DIR=$PWD/ray-tracing-c_b200/$b timeout 300 python tools/ab_render.py --frames 1:1200:300,6:600:300,7:1000:100 --reps 3 >> gpurun_out/r1_ab.jsonl 2>> gpurun_out/r1_ab.err
done
cat gpurun_out/r1_ab.jsonl
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mega_render_kernel -c 1 -o gpurun_out/r1_book1 -f python tools/ab_render.py --frames 1:1200:1000 --reps 0 > gpurun_out/r1_ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mega_render_kernel -c 1 -o gpurun_out/r1_book2 -f python tools/ab_render.py --frames 7:1000:100 --reps 0 > gpurun_out/r1_ncu2.log 2>&1
ls -la gpurun_out/
