# A/B helper: builds variants given as "file:sed-expr" pairs and benches each (240 steps)
# usage: bash tools/gpu_ab.sh  (variants are prepared as prebuilt .so copies under ab_variants/*.so)
for so in ab_variants/*.so; do
  cp $so marl-hideandseek_b200/lib/libhideseek_b200.so
  python bench.py --steps 240 --warmup 5 --no-cpu-baseline --profile-steps 1 > gpurun_out/ab.json 2>gpurun_out/ab.err || tail -3 gpurun_out/ab.err
  python -c "
import json; d=json.load(open('gpurun_out/ab.json')); print('$so', 'ms/step', round(d['ms_per_step'],4), 'value', round(d['value']))"
done
