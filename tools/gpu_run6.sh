#!/bin/bash
# 2 GPUs: the multi-GPU code paths (rt_render_multi with in-process NCCL, bench.py under torchrun with image_check and
# e2e_dropin), the full GPU suite on a 2-GPU box
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r6_gpus.txt
timeout 1800 python -m pytest tests -m gpu -q > gpurun_out/r6_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r6_pytest.log
tail -8 gpurun_out/r6_pytest.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r6_bench_n2.json 2> gpurun_out/r6_bench_n2.err; echo "bench2 rc=$?"
tail -c 3000 gpurun_out/r6_bench_n2.json; tail -5 gpurun_out/r6_bench_n2.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --config 4 --steps 3 --warmup 3 > gpurun_out/r6_bench_cfg4_n2.json 2> gpurun_out/r6_bench_cfg4_n2.err; echo "bench cfg4 rc=$?"
tail -c 1500 gpurun_out/r6_bench_cfg4_n2.json
B2="7:1000:100"; C="6:600:300"
ab() { fr=$1; shift; env "$@" timeout 300 python tools/ab_render.py --frames $fr --reps 3 --tag "$*" >> gpurun_out/r6_ab.jsonl 2>> gpurun_out/r6_ab.err; }
ab $C,$B2,5:1200:100 A=default
ab $B2 RTC_BUILD_DIR=$PWD/ray-tracing-c_b200/_build_m0
ab 5:1200:100 RT_MEGA_NV=0 RT_MEGA_MINB=8
python - <<'PY'
import json
for l in open('gpurun_out/r6_ab.jsonl'):
    d=json.loads(l); print(f"{d['tag'][-60:]:60s} s{d['scene']} {d['mseg_per_s']:9.1f}")
PY
