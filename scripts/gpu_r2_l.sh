#!/bin/bash
O=gpurun_out
V=$PWD/video_flow_ml_b200/variants
VFML_GEMM_NO_WS=1 python scripts/micro/ws_ab.py old > $O/l_old.txt 2>&1
VFML_WS_PAIR=0 python scripts/micro/ws_ab.py single-spin > $O/l_single.txt 2>&1
VFML_WS_PAIR=1 python scripts/micro/ws_ab.py pair-spin > $O/l_pair.txt 2>&1
VFML_WS_PAIR=1 VFML_LIB_PATH=$V/libvfml_nospin.so python scripts/micro/ws_ab.py pair-nospin > $O/l_pair_nospin.txt 2>&1
VFML_WS_PAIR=0 VFML_LIB_PATH=$V/libvfml_nospin.so python scripts/micro/ws_ab.py single-nospin > $O/l_single_nospin.txt 2>&1
paste -d'|' $O/l_old.txt <(cut -c15- $O/l_single.txt) <(cut -c15- $O/l_single_nospin.txt) <(cut -c15- $O/l_pair.txt) <(cut -c15- $O/l_pair_nospin.txt)
