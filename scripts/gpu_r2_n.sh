#!/bin/bash
# ncu --set full captures: tensor-core depthwise kernel, pair-mode cluster GEMM, general GEMM (same layer)
O=gpurun_out
python scripts/micro/dw_only.py 2 > $O/n_dw_plain.txt 2>&1 || exit 1
VFML_WS_PAIR=1 python scripts/micro/gemm_one_shape.py > $O/n_gemm_plain.txt 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:dwconv_tc -s 6 -c 1 -o $O/r2_dwconv_tc -f python scripts/micro/dw_only.py 2 > $O/n_ncu_dw.log 2>&1
VFML_WS_PAIR=1 ncu --set full --clock-control none --import-source on -k regex:gemm_ws -s 2 -c 1 -o $O/r2_gemm_ws_pair -f python scripts/micro/gemm_one_shape.py > $O/n_ncu_ws.log 2>&1
VFML_GEMM_NO_WS=1 ncu --set full --clock-control none --import-source on -k regex:gemm_tc -s 2 -c 1 -o $O/r2_gemm_tc -f python scripts/micro/gemm_one_shape.py > $O/n_ncu_tc.log 2>&1
tail -2 $O/n_ncu_dw.log $O/n_ncu_ws.log $O/n_ncu_tc.log
ls -la $O/*.ncu-rep
