#!/bin/bash
O=gpurun_out
for r in 1 2 3; do
  for cfg in "on" "nopf" "nopf_nopyr"; do
    case $cfg in
      on) E="";;
      nopf) E="VFML_PREFETCH=0";;
      nopf_nopyr) E="VFML_PREFETCH=0 VFML_PYRAMID_CACHE=0";;
    esac
    env $E python bench.py --no-extras > $O/s_${cfg}_$r.json 2>/dev/null
    python -c "import json; d=json.loads(open('$O/s_${cfg}_$r.json').read().strip().splitlines()[-1]); print('$cfg $r  steps %d ms %.2f value %.2f e2e %.2f' % (d['steps'], d['ms_per_step'], d['value'], d['e2e']['value']))"
  done
done
