#!/bin/bash
# single-GPU lines of the other BASELINE configs
mkdir -p gpurun_out
for c in C2 C3 C4 C5; do
  steps=8; [ $c = C5 ] && steps=4
  timeout 1200 python bench.py --config $c --steps $steps --warmup 3 > gpurun_out/r2e_bench_$c.json 2> gpurun_out/r2e_bench_$c.err; echo "$c rc=$?"
  python - <<PY
import json
try:
    d=json.load(open('gpurun_out/r2e_bench_$c.json'))
    print('$c', d['metric'], 'value', round(d['value'],3), 'ms/step', round(d['ms_per_step'],2), 'e2e', round(d['e2e']['value'],3), d['config'].get('encoder_passes'), 'launches', d['gpu_launches'])
except Exception as e:
    print('$c failed', e)
PY
  tail -3 gpurun_out/r2e_bench_$c.err
done
