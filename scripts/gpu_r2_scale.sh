#!/bin/bash
# usage: gpu_r2_scale.sh N "C1 C3 C5"
N=$1; shift
mkdir -p gpurun_out
for c in $1; do
  steps=20; [ $c = C5 ] && steps=4; [ $c = C3 ] && steps=8
  timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --config $c --steps $steps --warmup 3 --no-cpu-baseline > gpurun_out/r2_scale_${c}_n$N.json 2> gpurun_out/r2_scale_${c}_n$N.err; echo "$c N=$N rc=$?"
  python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/r2_scale_${c}_n$N.json') if l.startswith('{')][-1])
    print('$c N=$N', 'value', round(d['value'],3), 'ms/step', round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],3), d['scaling'], d['config']['frames_per_rank'], d['config'].get('encoder_passes'))
except Exception as e:
    print('$c failed', e)
PY
  grep -v "^W\|Warning\|warn" gpurun_out/r2_scale_${c}_n$N.err | tail -4
done
