#!/bin/bash
O=gpurun_out
timeout 300 python bench.py --steps 2 --warmup 3 --no-extras --passes 1 > $O/nl_plain.json 2>&1 || exit 1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:corr_lookup_tiled -s 150 -c 1 -o $O/r2_lookup_tiled -f python bench.py --steps 2 --warmup 3 --no-extras --passes 1 > $O/nl_ncu.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:dwconv_tc_kernel -s 150 -c 1 -o $O/r2_dwconv15 -f python bench.py --steps 2 --warmup 3 --no-extras --passes 1 > $O/nl_ncu2.log 2>&1
ls -la $O/r2_lookup_tiled.ncu-rep $O/r2_dwconv15.ncu-rep
