#!/bin/bash
O=gpurun_out
VFML_GEMM_NO_WS=1 python scripts/microbench.py > $O/k_micro_old.txt 2>&1
python scripts/microbench.py > $O/k_micro_auto.txt 2>&1
VFML_WS_PAIR=1 python scripts/microbench.py > $O/k_micro_pair.txt 2>&1
VFML_WS_PAIR=0 python scripts/microbench.py > $O/k_micro_single.txt 2>&1
echo "old | auto | pair | single"
paste -d'|' <(cut -c1-54 $O/k_micro_old.txt) <(cut -c44-54 $O/k_micro_auto.txt) <(cut -c44-54 $O/k_micro_pair.txt) <(cut -c44-54 $O/k_micro_single.txt)
VFML_GEMM_TRACE=1 python scripts/micro/gemm_trace.py > $O/k_trace_auto.txt 2>&1
grep "trace\]" $O/k_trace_auto.txt | cut -c1-340
