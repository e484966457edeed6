#!/bin/bash
# 8 GPUs, final round-2 code: the headline configuration at N = 2, 4, 8 and configs 3, 4 at N = 8
mkdir -p gpurun_out
nproc > gpurun_out/r33_box.txt
run() { cfg=$1; n=$2; shift 2
  timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29800 + cfg * 10 + n)) \
    bench.py --gpus $n --config $cfg "$@" >> gpurun_out/r33_scale_cfg$cfg.jsonl 2>> gpurun_out/r33_scale_cfg$cfg.err
  echo "cfg $cfg N $n rc=$?"; }
for n in 8 4 2; do run 2 $n --steps 5 --warmup 3 --no-cpu-baseline; done
run 3 8 --steps 5 --warmup 3 --no-cpu-baseline --no-dropin
run 4 8 --steps 3 --warmup 3 --no-cpu-baseline --no-dropin
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/r33_scale_cfg*.jsonl')):
    for l in open(f):
        if not l.startswith('{'): continue
        d = json.loads(l); ic = d.get('image_check') or {}; dr = d.get('e2e_dropin') or {}
        print(f, d['n_gpus'], round(d['value']), round(d['ms_per_step'], 3), 'kernel', round(d['roofline']['kernel_ms'],3), 'e2e', round(d['e2e']['value']), round(d['e2e']['ms_per_step'],3),
              'reduce_ms', round(d['reduce_ms'], 3), 'check', ic.get('max_abs_diff_u8'), 'dropin', round(dr.get('ms_per_step', 0), 3))
PY
