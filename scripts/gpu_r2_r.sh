#!/bin/bash
O=gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > $O/r_pytest.txt 2>&1; echo "rc=$?" >> $O/r_pytest.txt; tail -5 $O/r_pytest.txt | cut -c1-200
for r in 1 2; do
  VFML_PREFETCH=0 python bench.py --steps 20 --warmup 3 --no-extras > $O/r_bench_off_$r.json 2>$O/r_bench_off_$r.err
  python bench.py --steps 20 --warmup 3 --no-extras > $O/r_bench_on_$r.json 2>$O/r_bench_on_$r.err
done
for f in $O/r_bench_*.json; do echo $f; python -c "import sys,json; d=json.loads(open('$f').read().strip().splitlines()[-1]); print('  ms %.2f value %.2f e2e %.2f' % (d['ms_per_step'], d['value'], d['e2e']['value']))"; done
