#!/bin/bash
# Round-2 closing run on one B200: smoke, default bench line (C1 with extras), reference arm, launch list of the bench
# command (ncu, only after the plain run exited 0), the other configs.
O=gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/f_smoke.txt 2>&1; echo "smoke rc=$?"; tail -2 $O/f_smoke.txt
timeout 900 python bench.py > $O/f_bench_C1.json 2> $O/f_bench_C1.err; echo "bench rc=$?"; tail -c 1500 $O/f_bench_C1.json; tail -3 $O/f_bench_C1.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $O/f_bench_reference.json 2> $O/f_bench_reference.err; echo "reference rc=$?"; tail -c 600 $O/f_bench_reference.json
timeout 300 python bench.py --steps 3 --warmup 3 --no-extras > $O/f_bench_plain.json 2>&1 || exit 1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv --log-file $O/f_launches.csv python bench.py --steps 3 --warmup 3 --no-extras > $O/f_ncu.log 2>&1; echo "ncu rc=$?"
python scripts/summarize_launches.py $O/f_launches.csv > $O/f_launches_summary.txt 2>&1; head -32 $O/f_launches_summary.txt
for c in C2 C3 C4 C5; do
  timeout 900 python bench.py --config $c --steps 8 --warmup 3 > $O/f_bench_$c.json 2> $O/f_bench_$c.err; echo "$c rc=$?"
  python -c "import json; d=json.loads(open('$O/f_bench_$c.json').read().strip().splitlines()[-1]); print('  $c ms %.2f value %.3f e2e %.3f' % (d['ms_per_step'], d['value'], d['e2e']['value']))"
done
