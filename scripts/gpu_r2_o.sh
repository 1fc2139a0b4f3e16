#!/bin/bash
O=gpurun_out
V=$PWD/video_flow_ml_b200/variants
for cfg in "base" "nbuf2" "e12nbuf2" "e16nbuf2"; do
  if [ $cfg = base ]; then L=$PWD/video_flow_ml_b200/libvfml.so; else L=$V/libvfml_$cfg.so; fi
  VFML_WS_PAIR=1 VFML_LIB_PATH=$L python scripts/micro/ws_ab.py pair-$cfg > $O/o_pair_$cfg.txt 2>&1
  VFML_WS_PAIR=0 VFML_LIB_PATH=$L python scripts/micro/ws_ab.py sgl-$cfg > $O/o_sgl_$cfg.txt 2>&1
done
VFML_GEMM_NO_WS=1 python scripts/micro/ws_ab.py old > $O/o_old.txt 2>&1
paste -d'|' $O/o_old.txt <(cut -c29- $O/o_pair_base.txt) <(cut -c29- $O/o_pair_nbuf2.txt) <(cut -c29- $O/o_pair_e12nbuf2.txt) <(cut -c29- $O/o_pair_e16nbuf2.txt)
echo single:
paste -d'|' $O/o_old.txt <(cut -c29- $O/o_sgl_base.txt) <(cut -c29- $O/o_sgl_nbuf2.txt) <(cut -c29- $O/o_sgl_e12nbuf2.txt) <(cut -c29- $O/o_sgl_e16nbuf2.txt)
