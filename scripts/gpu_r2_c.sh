#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_ops.py -m gpu -q -x -k "gemm or peg or twins" > gpurun_out/r2c_pytest_gemm.txt 2>&1; echo "gemm tests rc=$?"; tail -3 gpurun_out/r2c_pytest_gemm.txt
timeout 600 python scripts/microbench.py > gpurun_out/r2c_microbench_new.txt 2>&1; echo "microbench new rc=$?"
VFML_LIB_PATH=$PWD/video_flow_ml_b200/variants/libvfml_epi_v1.so timeout 600 python scripts/microbench.py > gpurun_out/r2c_microbench_v1.txt 2>&1; echo "microbench v1 rc=$?"
paste <(cut -c1-64 gpurun_out/r2c_microbench_v1.txt) <(cut -c45-64 gpurun_out/r2c_microbench_new.txt) | head -48
VFML_GEMM_TRACE=1 timeout 300 python scripts/micro/gemm_trace.py > gpurun_out/r2c_gemm_trace.txt 2>&1
grep "epi=23\|epi=5 " gpurun_out/r2c_gemm_trace.txt | cut -c1-330 | head -12
timeout 1500 python -m pytest tests -m gpu -q -s > gpurun_out/r2c_pytest_gpu.txt 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2c_pytest_gpu.txt
grep -E "passed|failed|FAILED" gpurun_out/r2c_pytest_gpu.txt | tail -8
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/r2c_bench.json 2> gpurun_out/r2c_bench.err; echo "bench rc=$?"; python - <<'PY'
import json
d=json.load(open('gpurun_out/r2c_bench.json'))
print({k:d[k] for k in ('value','ms_per_step','gpu_launches')}, d['e2e']['value'], d['value_uncached']['value'], d['roofline']['frac'], d['roofline']['us_per_iteration'], d['clocks'])
PY
tail -3 gpurun_out/r2c_bench.err
