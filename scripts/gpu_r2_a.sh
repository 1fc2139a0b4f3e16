#!/bin/bash
# Round 2, first GPU call: full GPU test suite, parity table + ablation, default bench line, launch list.
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q -s > gpurun_out/r2a_pytest_gpu.txt 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/r2a_pytest_gpu.txt
tail -5 gpurun_out/r2a_pytest_gpu.txt
timeout 600 python scripts/fp16_ablation.py > gpurun_out/r2a_fp16_ablation.txt 2>&1; tail -12 gpurun_out/r2a_fp16_ablation.txt
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; echo "bench rc=$?"; tail -c 3000 gpurun_out/r2a_bench.json; tail -5 gpurun_out/r2a_bench.err
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2a_smoke.txt 2>&1; tail -2 gpurun_out/r2a_smoke.txt
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file gpurun_out/r2a_launches.csv python bench.py --steps 3 --warmup 3 --no-extras > gpurun_out/r2a_ncu.log 2>&1; echo "ncu rc=$?"
python scripts/summarize_launches.py gpurun_out/r2a_launches.csv > gpurun_out/r2a_launches_summary.txt 2>&1; head -30 gpurun_out/r2a_launches_summary.txt
