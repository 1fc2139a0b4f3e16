#!/bin/bash
# prefetch on by default: full GPU suite, every config
O=gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -4
for c in C1 C2 C3 C4 C5; do
  timeout 900 python bench.py --config $c --no-extras > $O/z_bench_$c.json 2> $O/z_bench_$c.err
  python -c "
import json; d=json.loads(open('$O/z_bench_$c.json').read().strip().splitlines()[-1]); print('$c', round(d['value'],3), round(d['e2e']['value'],3), d['config']['passes'])" || tail -5 $O/z_bench_$c.err
done
