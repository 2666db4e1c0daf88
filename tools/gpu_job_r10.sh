# r10 evidence set: GPU test suite, bench lines (both arms), BASELINE.md section-4 configurations,
# ncu launch list, ncu --set full capture of every stage kernel.  Every ncu pass runs only after the
# same command has exited 0 without ncu.
set -x
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r10_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r10_pytest_gpu.log
tail -3 gpurun_out/r10_pytest_gpu.log
python bench.py > gpurun_out/r10_bench_n1.json 2> gpurun_out/r10_bench_n1.err || exit 1
python bench.py --impl reference --steps 20 --warmup 3 > gpurun_out/r10_bench_reference.json 2> gpurun_out/r10_bench_reference.err
timeout 600 python tools/bench_configs.py > gpurun_out/r10_configs.json 2> gpurun_out/r10_configs.err
python bench.py --steps 160 --warmup 5 --no-cpu-baseline --profile-steps 1 > gpurun_out/r10_plain.json 2>/dev/null || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -s 4100 -c 132 --csv --log-file gpurun_out/r10_launches_ncu.csv \
    python bench.py --steps 160 --warmup 5 --no-cpu-baseline --profile-steps 1 > gpurun_out/r10_ncu_launch.log 2>&1
ncu --set full --import-source on --clock-control none \
    -k regex:"k_solve|k_narrowphase|k_rays|k_broadphase|k_observations|k_leaf_refit|k_rewards|k_substep|k_action" -s 2300 -c 24 -f \
    -o gpurun_out/r10_full python bench.py --steps 120 --warmup 5 --no-cpu-baseline --profile-steps 1 > gpurun_out/r10_ncu_full.log 2>&1
