#!/bin/bash
# per-op microbench of prebuilt library variants, side by side: scripts/micro/ab_micro.sh build/libA.so build/libB.so ...
i=0
cols=""
for lib in "$@"; do
  cp $lib video_flow_ml_b200/libvfml.so
  python scripts/microbench.py > gpurun_out/micro_$i.txt 2>&1
  if [ $i -gt 0 ]; then cols="$cols <(cut -c44-54 gpurun_out/micro_$i.txt)"; fi
  i=$((i+1))
done
echo "variants: $@"
eval paste -d"'|'" "<(cut -c1-54 gpurun_out/micro_0.txt)" $cols
