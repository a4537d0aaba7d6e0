timeout 400 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
for i in 1 2; do
timeout 250 python bench.py --steps 10 --warmup 3 --no-cpu > gpurun_out/sw.json 2>gpurun_out/sw.err; python -c "
import json
d=json.loads(open('gpurun_out/sw.json').read().strip().splitlines()[-1])
print(round(d['ms_per_step'],3),d['clocks']['reasons'],{k:v['ms'] for k,v in d['stages'].items() if 'mlp' in k or 'grad' in k}, d.get('extract_vol'))"
done
