#!/bin/bash
O=gpurun_out
timeout 400 python -m pytest tests/test_gpu_ops.py -m gpu -x -q -k "gemm" > $O/i_pytest.txt 2>&1; echo "rc=$?" >> $O/i_pytest.txt; tail -3 $O/i_pytest.txt
python scripts/microbench.py > $O/i_micro_ws.txt 2>&1
paste -d'|' <(cut -c1-54 $O/h_micro_old.txt) <(cut -c44-54 $O/i_micro_ws.txt)
VFML_GEMM_TRACE=1 python scripts/micro/gemm_trace.py > $O/i_trace_ws.txt 2>&1
grep "trace\]" $O/i_trace_ws.txt | cut -c1-330 | head -12
