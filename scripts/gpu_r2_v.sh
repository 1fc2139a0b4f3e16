#!/bin/bash
# fused volume epilogue: parity, A/B timing, engine tests, bench
O=gpurun_out
timeout 600 python -m pytest tests/test_gpu_ops.py -x -q -m gpu -k "fused or tiled or volume or no_alloc" 2>&1 | tail -5
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -2
timeout 300 python scripts/micro/corr_fused_ab.py > $O/v_corr_fused_ab.txt 2>&1; cat $O/v_corr_fused_ab.txt | tail -6
timeout 900 python -m pytest tests/test_gpu_e2e.py -x -q -m gpu 2>&1 | tail -3
timeout 600 python bench.py --no-extras > $O/v_bench_C1.json 2> $O/v_bench_C1.err; tail -c 900 $O/v_bench_C1.json
VFML_CORR_FUSED=0 timeout 600 python bench.py --no-extras > $O/v_bench_C1_unfused.json 2> $O/v_bench_C1_unfused.err; tail -c 900 $O/v_bench_C1_unfused.json
for i in 1 2 3; do VFML_BENCH_TRACE=1 timeout 300 python bench.py --no-extras > $O/v_trace_$i.json 2> $O/v_trace_$i.err; grep -h "resident\|e2e per" $O/v_trace_$i.err | cut -c1-400; done
