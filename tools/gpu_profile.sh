set -x
python tools/umma_bench.py > gpurun_out/umma_bench.log 2>&1
python bench.py --steps 3 --warmup 3 --no-cpu --no-extract > gpurun_out/b2.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r1.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-extract > gpurun_out/ncu_l.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:mlp_ -s 18 -c 6 -o gpurun_out/mlp_r1 -f python bench.py --steps 2 --warmup 3 --no-cpu --no-extract > gpurun_out/ncu_f.log 2>&1
ls -la gpurun_out
