#!/bin/bash
# cluster GEMM (gemm_ws.cu): microbench against the previous kernel, trace, epilogue-warp variant
O=gpurun_out
V=$PWD/video_flow_ml_b200/variants
VFML_GEMM_NO_WS=1 python scripts/microbench.py > $O/h_micro_old.txt 2>&1
python scripts/microbench.py > $O/h_micro_ws.txt 2>&1
VFML_LIB_PATH=$V/libvfml_ws12.so python scripts/microbench.py > $O/h_micro_ws12.txt 2>&1
paste -d'|' <(cut -c1-54 $O/h_micro_old.txt) <(cut -c44-54 $O/h_micro_ws.txt) <(cut -c44-54 $O/h_micro_ws12.txt)
VFML_GEMM_TRACE=1 python scripts/micro/gemm_trace.py > $O/h_trace_ws.txt 2>&1
grep "trace\]" $O/h_trace_ws.txt | cut -c1-330
