#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s > gpurun_out/r2b_pytest_gpu.txt 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2b_pytest_gpu.txt
grep -E "EPE|tile|passed|failed|FAILED|Error" gpurun_out/r2b_pytest_gpu.txt | tail -40
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2b_smoke.txt 2>&1; tail -2 gpurun_out/r2b_smoke.txt
timeout 600 python scripts/diag_cached.py > gpurun_out/r2b_diag_cached.txt 2>&1; cat gpurun_out/r2b_diag_cached.txt | tail -8
VFML_GEMM_TRACE=1 timeout 300 python scripts/micro/gemm_trace.py > gpurun_out/r2b_gemm_trace.txt 2>&1; grep -c trace gpurun_out/r2b_gemm_trace.txt
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest tests/test_gpu_ops.py -m gpu -q -x -k "gemm_tensor_core or dwconv_tc or tiled" > gpurun_out/r2b_memcheck.txt 2>&1; echo "memcheck rc=$?"; tail -4 gpurun_out/r2b_memcheck.txt
timeout 900 compute-sanitizer --tool racecheck --error-exitcode 9 python -m pytest tests/test_gpu_ops.py -m gpu -q -x -k "dwconv_tc or tiled" > gpurun_out/r2b_racecheck.txt 2>&1; echo "racecheck rc=$?"; tail -4 gpurun_out/r2b_racecheck.txt
