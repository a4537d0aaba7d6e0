#!/bin/bash
# Round profile capture (run through gpurun): bench line, launch list, ncu --set full of the MLP kernels of one step.
set -x
python bench.py --steps 20 --warmup 5 > gpurun_out/bench_full.json 2> gpurun_out/bench_full.err || exit 1
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-extract > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:mlp_ -s 18 -c 6 -o gpurun_out/mlp_full -f python bench.py --steps 2 --warmup 3 --no-cpu --no-extract > gpurun_out/ncu_f.log 2>&1
python tools/umma_bench.py > gpurun_out/umma_bench.log 2>&1
python tools/store_bench.py > gpurun_out/store_bench.log 2>&1
python tools/pair_mma_bench.py > gpurun_out/pair_mma_bench.log 2>&1
ls -la gpurun_out
