#!/bin/bash
# Round-3 profile capture (run through gpurun): launch list of a train step and ncu --set full of the MLP kernels of one step.
# Every ncu command runs only after the same command exited 0 without ncu.
set -x
python bench.py --steps 2 --warmup 3 --no-cpu --no-extract --min-seconds 0.01 > gpurun_out/plain_bench.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r3_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-extract --min-seconds 0.01 > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:mlp_ -s 18 -c 6 -o gpurun_out/r3_mlp_full -f python bench.py --steps 2 --warmup 3 --no-cpu --no-extract --min-seconds 0.01 > gpurun_out/ncu_f.log 2>&1
ls -la gpurun_out | tail -5
