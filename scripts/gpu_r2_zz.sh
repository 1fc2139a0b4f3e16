#!/bin/bash
# pyramids beside the attention (side stream): full GPU suite with it on, bench on / off
O=gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
for v in 1 0 1 0; do
  VFML_PYRAMID_SIDE=$v timeout 300 python bench.py --no-extras --no-cpu-baseline > $O/zz_side_$v.json 2> $O/zz_side_$v.err
  python -c "
import json; d=json.loads(open('$O/zz_side_$v.json').read().strip().splitlines()[-1]); print('side=$v', round(d['value'],3), round(d['e2e']['value'],3), d['config']['passes'])" || tail -3 $O/zz_side_$v.err
done
