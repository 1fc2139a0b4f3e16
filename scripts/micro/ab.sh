#!/bin/bash
# interleaved A/B of prebuilt library variants: scripts/micro/ab.sh build/libA.so build/libB.so ...
for round in 1 2; do
  for lib in "$@"; do
    cp $lib video_flow_ml_b200/libvfml.so
    echo "== $lib (round $round)  sm clock: $(nvidia-smi --query-gpu=clocks.sm --format=csv,noheader)"
    python scripts/micro/gemm_probe.py > gpurun_out/ab.txt 2>&1; head -5 gpurun_out/ab.txt | cut -c1-140
  done
done
