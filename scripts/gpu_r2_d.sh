#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s > gpurun_out/r2d_pytest_gpu.txt 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2d_pytest_gpu.txt
grep -E "passed|failed|FAILED" gpurun_out/r2d_pytest_gpu.txt | tail -8
timeout 600 python scripts/microbench.py > gpurun_out/r2d_microbench_main.txt 2>&1
VFML_LIB_PATH=$PWD/video_flow_ml_b200/variants/libvfml_epi_v2.so timeout 600 python scripts/microbench.py > gpurun_out/r2d_microbench_v2.txt 2>&1
VFML_LIB_PATH=$PWD/video_flow_ml_b200/variants/libvfml_epi16.so timeout 600 python scripts/microbench.py > gpurun_out/r2d_microbench_epi16.txt 2>&1
paste <(cut -c1-64 gpurun_out/r2d_microbench_main.txt) <(cut -c45-64 gpurun_out/r2d_microbench_v2.txt) <(cut -c45-64 gpurun_out/r2d_microbench_epi16.txt) | head -48
VFML_GEMM_TRACE=1 timeout 300 python scripts/micro/gemm_trace.py > gpurun_out/r2d_gemm_trace_v1.txt 2>&1
VFML_LIB_PATH=$PWD/video_flow_ml_b200/variants/libvfml_epi_v2.so VFML_GEMM_TRACE=1 timeout 300 python scripts/micro/gemm_trace.py > gpurun_out/r2d_gemm_trace_v2.txt 2>&1
grep "epi=23\|epi=5 " gpurun_out/r2d_gemm_trace_v1.txt | cut -c1-330 | head -6
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/r2d_bench.json 2> gpurun_out/r2d_bench.err; echo "bench rc=$?"; python - <<'PY'
import json
d=json.load(open('gpurun_out/r2d_bench.json'))
print({k:d[k] for k in ('value','ms_per_step','gpu_launches')}, 'e2e', d['e2e']['value'], 'uncached', d['value_uncached']['value'], 'gemm frac', d['roofline']['frac'], d['roofline']['us_per_iteration'], d['clocks'])
PY
tail -3 gpurun_out/r2d_bench.err
