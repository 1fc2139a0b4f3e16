#!/bin/bash
# N = 1, 2, 4, 8 back to back on one 8-GPU box (gpurun --gpus 8): the driver's SCALE run, here for the record.
mkdir -p gpurun_out
python bench.py --gpus 1 --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/scale_1.json 2> gpurun_out/scale_1.err
for n in 2 4 8; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) \
      bench.py --gpus $n --steps 8 --warmup 3 --no-cpu-baseline > gpurun_out/scale_$n.json 2> gpurun_out/scale_$n.err
done
python - <<'PY'
import json
base = None
for n in (1, 2, 4, 8):
    try:
        d = json.loads(open(f"gpurun_out/scale_{n}.json").read().strip().splitlines()[-1])
    except Exception as e:
        print(n, "failed", e); continue
    base = base or d["value"]
    print(f"N={n}: value {d['value']:.2f} frames/s ({d['value'] / base:.2f}x), e2e {d['e2e']['value']:.2f}, ms_per_step {d['ms_per_step']:.2f}")
PY
