#!/bin/bash
# depthwise kernel variants side by side (run under gpurun).  Variant libraries (prebuilt, scripts/micro/libvfml_<v>.so):
# build the variants first:  VFML_NVCC_EXTRA="-DVFML_DW_K8=0" VFML_LIB_OUT=scripts/micro/libvfml_u.so python -m video_flow_ml_b200.build
#                             VFML_NVCC_EXTRA="-DVFML_DW_K8=1" VFML_LIB_OUT=scripts/micro/libvfml_u_k8.so python -m video_flow_ml_b200.build
#   u = baseline MMA loop, u_k8 = k16 + k8 MMA split; VFML_DW_TMA: bit 0 TMA load, bit 1 TMA store; VFML_DW_TH: tile rows
mkdir -p gpurun_out
cp video_flow_ml_b200/libvfml.so /tmp/libvfml_orig.so
for v in u u_k8; do
  cp scripts/micro/libvfml_$v.so video_flow_ml_b200/libvfml.so
  for th in 64 32; do
  for m in 0 3; do
    echo "== variant $v TH=$th VFML_DW_TMA=$m: tests"; VFML_DW_TH=$th VFML_DW_TMA=$m timeout 200 python -m pytest tests/test_gpu_ops.py -m gpu -x -q -k dwconv 2>&1 | tail -2
  done
  for m in 0 1 2 3; do echo "== variant $v TH=$th VFML_DW_TMA=$m: timing"; VFML_DW_TH=$th VFML_DW_TMA=$m timeout 100 python scripts/micro/dw_only.py 2>&1 | tail -5; done
  done
done
cp /tmp/libvfml_orig.so video_flow_ml_b200/libvfml.so
