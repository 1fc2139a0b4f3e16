#!/bin/bash
# depthwise kernel variants side by side (run under gpurun): paired / unpaired MMA loop x TMA modes 0-3
mkdir -p gpurun_out
echo "== dwconv tests, default (mode 0, paired)"; timeout 200 python -m pytest tests/test_gpu_ops.py -m gpu -x -q -k dwconv 2>&1 | tail -2
for m in 1 2 3; do echo "== dwconv tests, VFML_DW_TMA=$m"; VFML_DW_TMA=$m timeout 200 python -m pytest tests/test_gpu_ops.py -m gpu -x -q -k dwconv 2>&1 | tail -2; done
for m in 0 1 2 3; do echo "== timing paired, VFML_DW_TMA=$m"; VFML_DW_TMA=$m timeout 100 python scripts/micro/dw_only.py; done
cp video_flow_ml_b200/libvfml.so /tmp/libvfml_paired.so
cp scripts/micro/libvfml_unpaired.so video_flow_ml_b200/libvfml.so
for m in 0 3; do echo "== timing unpaired, VFML_DW_TMA=$m"; VFML_DW_TMA=$m timeout 100 python scripts/micro/dw_only.py; done
cp /tmp/libvfml_paired.so video_flow_ml_b200/libvfml.so
