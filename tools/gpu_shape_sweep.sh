# step-graph shape sweep: HIDESEEK_STEP_SLICES x HIDESEEK_STEP_BRANCHES, 240 steps each
for sb in "1 1" "2 1" "2 2" "3 3" "4 2" "4 4" "6 3" "8 4"; do
  set -- $sb
  HIDESEEK_STEP_SLICES=$1 HIDESEEK_STEP_BRANCHES=$2 timeout 120 python bench.py --steps 240 --warmup 5 --no-cpu-baseline --profile-steps 1 > gpurun_out/shape.json 2>gpurun_out/shape.err || { echo "$sb FAILED"; continue; }
  python -c "
import json; d=json.load(open('gpurun_out/shape.json')); print('slices/branches $1/$2', 'ms/step', round(d['ms_per_step'],4), 'full_episode', round(d['full_episode']['ms_per_step'],4))"
done
