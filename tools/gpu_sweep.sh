# throughput vs worlds per GPU (the metric is quoted at 16000; the sweep shows where the GPU saturates)
for n in 4000 8000 16000 32000 64000 128000; do
  python bench.py --worlds $n --steps 240 --warmup 5 --no-cpu-baseline --profile-steps 1 > gpurun_out/sweep_$n.json 2>gpurun_out/sweep.err || tail -3 gpurun_out/sweep.err
  python -c "
import json; d=json.load(open('gpurun_out/sweep_$n.json')); print('worlds $n ms/step', round(d['ms_per_step'],4), 'env-steps/s', round(d['value']), 'e2e', round(d['e2e']['value']))"
done
