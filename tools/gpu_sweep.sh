for cfg in "1 1" "2 2" "3 3" "4 4" "4 2"; do
  set -- $cfg
  HIDESEEK_STEP_SLICES=$1 HIDESEEK_STEP_BRANCHES=$2 python bench.py --steps 240 --warmup 5 --no-cpu-baseline --profile-steps 1 > gpurun_out/sweep.json
  python -c "
import json; d=json.load(open('gpurun_out/sweep.json')); print('slices $1 branches $2 ms/step', round(d['ms_per_step'],4), 'value', round(d['value']), 'e2e', round(d['e2e']['value']))"
done
