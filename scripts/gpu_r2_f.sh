#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_ops.py -m gpu -q -x -k "attention or twins or peg" > gpurun_out/r2f_pytest_attn.txt 2>&1; echo "attn tests rc=$?"; tail -4 gpurun_out/r2f_pytest_attn.txt
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r2f_pytest_gpu.txt 2>&1; echo "pytest rc=$?"; grep -E "passed|failed|FAILED" gpurun_out/r2f_pytest_gpu.txt | tail -8
timeout 300 python scripts/diag_cached2.py 2>&1 | tail -9
timeout 600 python bench.py --config C4 --steps 16 --warmup 3 > gpurun_out/r2f_bench_C4.json 2> gpurun_out/r2f_bench_C4.err; echo "C4 rc=$?"; tail -c 600 gpurun_out/r2f_bench_C4.json; tail -3 gpurun_out/r2f_bench_C4.err
