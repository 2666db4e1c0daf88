python bench.py --steps 240 --warmup 5 --no-cpu-baseline --profile-steps 8 > gpurun_out/vote.json && python -c "
import json; d=json.load(open('gpurun_out/vote.json')); print('base ms/step', round(d['ms_per_step'],4), [ (k['kernel'], round(k['avg_ms'],4)) for k in d['kernels'] if k['kernel'] in ('k_rays','k_rewards')])"
cd marl-hideandseek_b200/lib
cp libhideseek_b200.so /tmp/orig.so
cp libhs_ordered.so libhideseek_b200.so
cd ../..
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 240 --warmup 5 --no-cpu-baseline --profile-steps 8 > gpurun_out/vote.json && python -c "
import json; d=json.load(open('gpurun_out/vote.json')); print('ordered ms/step', round(d['ms_per_step'],4), [ (k['kernel'], round(k['avg_ms'],4)) for k in d['kernels'] if k['kernel'] in ('k_rays','k_rewards')])"
cp /tmp/orig.so marl-hideandseek_b200/lib/libhideseek_b200.so
