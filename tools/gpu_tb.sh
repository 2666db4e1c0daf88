python -m pytest tests -m gpu -x -q 2>&1 | tail -6; python bench.py --steps 240 --warmup 5 --no-cpu-baseline > gpurun_out/cur_bench.json; python -c "
import json; d=json.load(open('gpurun_out/cur_bench.json')); print(d['ms_per_step'], d['value'], 'e2e', d['e2e']['value']); [print(k['kernel'], round(k['avg_ms'],4), k['launches_per_step']) for k in d['kernels']]"
