# r06 evidence set: bench lines (both arms), ncu launch list, ncu --set full capture of every stage kernel.
# Every ncu pass runs only after the same command has exited 0 without ncu.
set -x
python bench.py > gpurun_out/r06_bench_n1.json 2> gpurun_out/r06_bench_n1.err || exit 1
python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/r06_bench_reference.json 2> gpurun_out/r06_bench_reference.err
python bench.py --steps 160 --warmup 5 --no-cpu-baseline --profile-steps 1 > gpurun_out/r06_plain.json 2>/dev/null || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -s 4800 -c 150 --csv --log-file gpurun_out/r06_launches_ncu.csv \
    python bench.py --steps 160 --warmup 5 --no-cpu-baseline --profile-steps 1 > gpurun_out/r06_ncu_launch.log 2>&1
ncu --set full --import-source on --clock-control none \
    -k regex:"k_solve|k_narrowphase|k_rays|k_broadphase|k_observations|k_leaf_refit|k_rewards|k_substep|k_action" -s 2700 -c 28 -f \
    -o gpurun_out/r06_full python bench.py --steps 120 --warmup 5 --no-cpu-baseline --profile-steps 1 > gpurun_out/r06_ncu_full.log 2>&1
