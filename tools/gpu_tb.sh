#!/bin/bash
# pytest -m gpu, then a 240-step bench with the per-kernel table
python -m pytest tests -x -q -m gpu 2>&1 | tail -5
python bench.py --steps 240 --warmup 5 --no-cpu-baseline --profile-steps 8 > gpurun_out/tb.json 2> gpurun_out/tb.err || tail -5 gpurun_out/tb.err
python - <<'PY'
import json
d = json.load(open('gpurun_out/tb.json'))
print('ms/step', round(d['ms_per_step'], 4), 'value', round(d['value']), 'e2e', round(d['e2e']['value']))
for k in d.get('kernels', [])[:12]:
    print(k)
PY
