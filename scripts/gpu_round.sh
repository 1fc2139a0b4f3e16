#!/bin/bash
# One gpurun call: GPU tests, smoke, parity report, bench, microbench. Outputs under gpurun_out/.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke.log
python scripts/parity_report.py > gpurun_out/parity.log 2>&1; cat gpurun_out/parity.log | tail -8
python bench.py --steps 8 --warmup 3 > gpurun_out/bench_fp16.json 2> gpurun_out/bench_fp16.err; echo "bench fp16 rc=$?"; cat gpurun_out/bench_fp16.json; tail -3 gpurun_out/bench_fp16.err
if [ "$1" = "full" ]; then
python bench.py --precision fp32 --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/bench_fp32.json 2> gpurun_out/bench_fp32.err; echo "bench fp32 rc=$?"; cat gpurun_out/bench_fp32.json
python scripts/microbench.py > gpurun_out/microbench.log 2>&1; echo "microbench rc=$?"; cat gpurun_out/microbench.log
fi
