#!/bin/bash
# host-buffer call overhead breakdown; ncu capture of the fused volume kernel (after the same command ran plain)
O=gpurun_out
timeout 500 python scripts/micro/e2e_tail.py > $O/x_e2e_tail.txt 2>&1; tail -12 $O/x_e2e_tail.txt
timeout 300 python scripts/micro/corr_fused_ab.py > $O/x_ab.txt 2>&1 || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:corr_fused_kernel -s 4 -c 1 -o $O/r2_corr_fused -f python scripts/micro/corr_fused_ab.py > $O/x_ncu.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file $O/x_launches.csv python scripts/micro/corr_fused_ab.py > $O/x_ncu2.log 2>&1
ls -la $O/r2_corr_fused.ncu-rep; head -30 $O/x_launches.csv | cut -c1-200
