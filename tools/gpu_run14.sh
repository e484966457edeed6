#!/bin/bash
# 2 GPUs, final build: the full GPU suite (incl. the 2-GPU tests) and bench.py at N = 2 (cached DistRenderer e2e path)
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r14_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r14_pytest.log
tail -5 gpurun_out/r14_pytest.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/r14_bench_cfg2_n2.json 2> gpurun_out/r14_bench_cfg2_n2.err; echo "bench cfg2 N=2 rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29712 bench.py --gpus 2 --config 3 --steps 5 --warmup 3 > gpurun_out/r14_bench_cfg3_n2.json 2> gpurun_out/r14_bench_cfg3_n2.err; echo "bench cfg3 N=2 rc=$?"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29713 bench.py --impl reference --gpus 2 --steps 1 --warmup 0 --ref-spp 8 > gpurun_out/r14_ref_n2.json 2> gpurun_out/r14_ref_n2.err; echo "ref N=2 rc=$?"; wc -l gpurun_out/r14_ref_n2.json
python - <<'PY'
import json
for f in ('gpurun_out/r14_bench_cfg2_n2.json','gpurun_out/r14_bench_cfg3_n2.json'):
    try:
        d=json.loads([l for l in open(f) if l.startswith('{')][-1])
        print(f, d['n_gpus'], round(d['value']), round(d['ms_per_step'],3), 'kernel', round(d['roofline']['kernel_ms'],3), 'e2e ms', round(d['e2e']['ms_per_step'],3), d['e2e']['ms_steps'], 'reduce', round(d['reduce_ms'],3), 'chk', d['image_check']['max_abs_diff_u8'], 'dropin', (d.get('e2e_dropin') or {}).get('ms_per_step'), 'exec_frac', d['roofline'].get('executed_frac'), 'launches', d['gpu_launches'])
    except Exception as e:
        print(f, 'ERR', e); print(open(f.replace('.json','.err')).read()[-1500:])
PY
