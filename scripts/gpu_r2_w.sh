#!/bin/bash
# feature-cache arena: tests, then the default bench three times with the per-call host trace (are the cudaMalloc stalls gone?)
O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_e2e.py -x -q -m gpu 2>&1 | tail -3
for i in 1 2 3; do VFML_BENCH_TRACE=1 timeout 300 python bench.py --no-extras > $O/w_trace_$i.json 2> $O/w_trace_$i.err; grep -h "resident pass" $O/w_trace_$i.err | cut -c1-200; python -c "
import json; d=json.loads(open('$O/w_trace_$i.json').read().strip().splitlines()[-1]); print(round(d['value'],2), round(d['e2e']['value'],2), d['config']['passes'])"; done
for c in C2 C3 C5; do timeout 600 python bench.py --config $c --no-extras > $O/w_bench_$c.json 2> $O/w_bench_$c.err; python -c "
import json; d=json.loads(open('$O/w_bench_$c.json').read().strip().splitlines()[-1]); print('$c', round(d['value'],3), round(d['e2e']['value'],3), d['config']['passes'])"; done
