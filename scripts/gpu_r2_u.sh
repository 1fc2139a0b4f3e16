#!/bin/bash
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $O/u_pytest.txt 2>&1; echo "rc=$?" >> $O/u_pytest.txt; tail -4 $O/u_pytest.txt | cut -c1-200
for r in 1 2; do
  VFML_GEMM_NO_AUTOTUNE=1 python bench.py --steps 20 --warmup 3 --no-extras > $O/u_bench_off_$r.json 2>$O/u_bench_off_$r.err
  VFML_GEMM_AUTOTUNE_LOG=1 python bench.py --steps 20 --warmup 3 --no-extras > $O/u_bench_on_$r.json 2>$O/u_bench_on_$r.err
done
grep "autotune\]" $O/u_bench_on_1.err | sort -u | head -60
for f in $O/u_bench_*.json; do echo $f; python -c "import sys,json; d=json.loads(open('$f').read().strip().splitlines()[-1]); print('  ms %.2f value %.2f e2e %.2f' % (d['ms_per_step'], d['value'], d['e2e']['value']))"; done
