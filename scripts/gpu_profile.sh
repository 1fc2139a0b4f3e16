#!/bin/bash
# ncu evidence for the bench command (run under gpurun): per-launch time list + one full capture of the roofline kernel.
mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_plain.json 2> gpurun_out/bench_plain.err || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:corr_lookup_tiled -s 200 -c 1 -o gpurun_out/lookup_full -f \
    python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_lookup.log 2>&1
tail -2 gpurun_out/ncu_lookup.log
