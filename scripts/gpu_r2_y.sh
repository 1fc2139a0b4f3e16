#!/bin/bash
# fused volume for C = 256 (the engine's feature maps) + ring order of the per-frame state: tests, A/B, bench
O=gpurun_out
timeout 600 python -m pytest tests/test_gpu_ops.py -x -q -m gpu -k "fused or tiled or volume or misc_update" 2>&1 | tail -4
timeout 900 python -m pytest tests/test_gpu_e2e.py tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -4
timeout 300 python scripts/micro/corr_fused_ab.py > $O/y_corr_fused_ab.txt 2>&1; tail -6 $O/y_corr_fused_ab.txt
run() { # name, env...
  name=$1; shift
  env "$@" timeout 600 python bench.py --no-extras > $O/y_bench_$name.json 2> $O/y_bench_$name.err
  python -c "
import json; d=json.loads(open('$O/y_bench_$name.json').read().strip().splitlines()[-1]); print('$name', round(d['value'],3), round(d['e2e']['value'],3), d['ms_per_step'], d['config']['passes'], d['gpu_launches'])" || tail -5 $O/y_bench_$name.err
}
run ring_fused X=1
run ring_unfused VFML_CORR_FUSED=0
run copy_fused VFML_PYRAMID_RING=0
run copy_unfused VFML_PYRAMID_RING=0 VFML_CORR_FUSED=0
run ring_fused_prefetch VFML_PREFETCH=1
run ring_fused_2 X=1
env timeout 600 python bench.py --config C3 --no-extras > $O/y_bench_C3.json 2> $O/y_bench_C3.err; python -c "
import json; d=json.loads(open('$O/y_bench_C3.json').read().strip().splitlines()[-1]); print('C3', round(d['value'],3), round(d['e2e']['value'],3), d['config']['passes'])"
