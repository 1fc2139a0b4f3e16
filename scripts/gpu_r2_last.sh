#!/bin/bash
# cold windows through one batched encoder pass again (VFML_ENCODE_PER_FRAME=0), now that the feature cache does not allocate: bench + full suite
O=gpurun_out
for v in 0 1; do
  VFML_ENCODE_PER_FRAME=$v timeout 300 python bench.py --no-extras --no-cpu-baseline > $O/last_epf_$v.json 2> $O/last_epf_$v.err
  python -c "
import json; d=json.loads(open('$O/last_epf_$v.json').read().strip().splitlines()[-1]); print('encode_per_frame=$v', round(d['value'],3), round(d['e2e']['value'],3), d['config']['passes'])" || tail -3 $O/last_epf_$v.err
done
VFML_ENCODE_PER_FRAME=0 timeout 600 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
