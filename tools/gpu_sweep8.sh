#!/bin/bash
# 8 GPUs: every BASELINE.json configuration at N = 2, 4, 8 (N = 1 comes from the 1-GPU runs), config 5 included
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r7_gpus.txt; nproc >> gpurun_out/r7_gpus.txt
run() { # config N extra...
  cfg=$1; n=$2; shift 2
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600 + cfg * 10 + n)) \
    bench.py --gpus $n --config $cfg "$@" >> gpurun_out/r7_scale_cfg$cfg.jsonl 2>> gpurun_out/r7_scale_cfg$cfg.err
  echo "cfg $cfg N $n rc=$?"
}
for n in 2 4 8; do run 2 $n --steps 5 --warmup 3 --no-cpu-baseline; done
for n in 2 4 8; do run 1 $n --steps 5 --warmup 3 --no-cpu-baseline --no-dropin; done
for n in 2 4 8; do run 3 $n --steps 5 --warmup 3 --no-cpu-baseline --no-dropin; done
for n in 2 4 8; do run 4 $n --steps 3 --warmup 3 --no-cpu-baseline --no-dropin; done
run 5 8 --steps 1 --warmup 3 --warmup-spp 64 --no-cpu-baseline
run 5 4 --steps 1 --warmup 3 --warmup-spp 64 --no-cpu-baseline --no-dropin --no-image-check
run 5 2 --steps 1 --warmup 3 --warmup-spp 64 --no-cpu-baseline --no-dropin --no-image-check
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/r7_scale_cfg*.jsonl')):
    for l in open(f):
        if not l.startswith('{'): continue
        d = json.loads(l)
        ic = d.get('image_check') or {}
        dr = d.get('e2e_dropin') or {}
        print(f, d['n_gpus'], round(d['value']), round(d['ms_per_step'], 2), 'e2e', round(d['e2e']['value']), 'reduce_ms', round(d['reduce_ms'], 3),
              'check', ic.get('max_abs_diff_u8'), 'dropin', round(dr.get('ms_per_step', 0), 2))
PY
