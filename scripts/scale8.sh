#!/bin/bash
# N = 8 (and 2) on one 8-GPU box: validates the max-over-ranks timing and the gather inside the timed region.
for n in 8 2; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
      bench.py --gpus $n --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/scale_$n.json 2> gpurun_out/scale_$n.err
  python -c "
import json; d = json.loads(open('gpurun_out/scale_$n.json').read().strip().splitlines()[-1])
print('N=$n: value %.2f frames/s, e2e %.2f, ms_per_step %.2f' % (d['value'], d['e2e']['value'], d['ms_per_step']))"
done
