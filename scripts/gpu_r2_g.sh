#!/bin/bash
# A/B of the tile-ahead residual prefetch (variants/libvfml_resahead.so) against the committed epilogue (libvfml.so)
V=$PWD/video_flow_ml_b200/variants/libvfml_resahead.so
O=gpurun_out
VFML_LIB_PATH=$V python -m pytest tests/test_gpu_ops.py tests/test_gpu_e2e.py -m gpu -x -q > $O/g_pytest_variant.txt 2>&1; echo "variant pytest rc=$?" >> $O/g_pytest_variant.txt
python scripts/microbench.py > $O/g_micro_base.txt 2>&1
VFML_LIB_PATH=$V python scripts/microbench.py > $O/g_micro_var.txt 2>&1
for r in 1 2; do
  python bench.py --steps 20 --warmup 3 --no-extras > $O/g_bench_base_$r.json 2>$O/g_bench_base_$r.err
  VFML_LIB_PATH=$V python bench.py --steps 20 --warmup 3 --no-extras > $O/g_bench_var_$r.json 2>$O/g_bench_var_$r.err
done
VFML_GEMM_TRACE=1 python scripts/micro/gemm_trace.py > $O/g_trace_base.txt 2>&1
VFML_GEMM_TRACE=1 VFML_LIB_PATH=$V python scripts/micro/gemm_trace.py > $O/g_trace_var.txt 2>&1
tail -n 3 $O/g_pytest_variant.txt
paste -d'|' <(cut -c1-60 $O/g_micro_base.txt) <(cut -c44-60 $O/g_micro_var.txt)
for f in $O/g_bench_*.json; do echo $f; python -c "import sys,json; d=json.loads(open('$f').read().strip().splitlines()[-1]); print('  ms %.2f value %.2f e2e %.2f' % (d['ms_per_step'], d['value'], d['e2e']['value']))"; done
