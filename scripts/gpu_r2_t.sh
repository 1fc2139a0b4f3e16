#!/bin/bash
O=gpurun_out
timeout 400 python -m pytest tests/test_gpu_ops.py -m gpu -x -q -k "gemm" > $O/t_pytest.txt 2>&1; echo "rc=$?" >> $O/t_pytest.txt; tail -4 $O/t_pytest.txt | cut -c1-200
VFML_GEMM_NO_WS=1 python scripts/micro/ws_ab.py old > $O/t_old.txt 2>&1
for pc in "1 4" "1 6" "1 8" "0 3" "0 4" "0 6"; do set -- $pc
  VFML_WS_PAIR=$1 VFML_WS_CS=$2 python scripts/micro/ws_ab.py p$1cs$2 > $O/t_p$1_cs$2.txt 2>&1
done
paste -d'|' <(cut -c16-50 $O/t_old.txt) <(cut -c30-50 $O/t_p1_cs4.txt) <(cut -c30-50 $O/t_p1_cs6.txt) <(cut -c30-50 $O/t_p1_cs8.txt) <(cut -c30-50 $O/t_p0_cs3.txt) <(cut -c30-50 $O/t_p0_cs4.txt) <(cut -c30-50 $O/t_p0_cs6.txt)
