# A/B helper: benches every prebuilt library variant under ab_variants/*.so (240 steps each),
# restoring the in-tree library afterwards.  Every run is bounded by `timeout`.
keep=/tmp/hs_keep_$$.so
cp marl-hideandseek_b200/lib/libhideseek_b200.so $keep
for so in ab_variants/*.so; do
  cp $so marl-hideandseek_b200/lib/libhideseek_b200.so
  timeout 120 python bench.py --steps ${AB_STEPS:-240} --warmup 5 --no-cpu-baseline --profile-steps 2 > gpurun_out/ab.json 2>gpurun_out/ab.err || { echo "$so FAILED"; tail -3 gpurun_out/ab.err; continue; }
  python -c "
import json; d=json.load(open('gpurun_out/ab.json')); k={x['kernel']:x['avg_ms'] for x in d['kernels']}
print('$so', 'ms/step', round(d['ms_per_step'],4), 'value', round(d['value']), ' '.join(f'{n[2:]}={k[n]:.4f}' for n in list(k)[:4]))"
done
cp $keep marl-hideandseek_b200/lib/libhideseek_b200.so
