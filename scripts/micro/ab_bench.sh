#!/bin/bash
# interleaved end-to-end A/B of prebuilt library variants: scripts/micro/ab_bench.sh build/libA.so build/libB.so ...
for round in 1 2; do
  for lib in "$@"; do
    cp $lib video_flow_ml_b200/libvfml.so
    echo "== $lib (round $round)"
    python bench.py --steps 6 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('  ms_per_step %.2f  value %.2f  e2e %.2f' % (d['ms_per_step'], d['value'], d['e2e']['value']))"
  done
done
for lib in "$@"; do
  cp $lib video_flow_ml_b200/libvfml.so
  echo "== parity $lib"
  python scripts/parity_report.py 2>&1 | grep "gain 1.0 tier fp16"
done
