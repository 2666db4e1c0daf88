ncu --set full --import-source on --clock-control none -k regex:"^k_solve$|k_narrowphase|k_rays" -s 900 -c 5 -f -o gpurun_out/r06_hot python bench.py --steps 120 --warmup 5 --no-cpu-baseline --profile-steps 1 > gpurun_out/r06_ncu_hot.log 2>&1; echo ncu rc=$?
timeout 600 python tools/cpu_config0.py --out gpurun_out/r06_config0.json > gpurun_out/r06_config0.log 2>&1; echo config0 rc=$?
for tool in memcheck racecheck initcheck; do
  STEPS=6 timeout 420 compute-sanitizer --tool $tool --kernel-regex kns=hs:: --print-limit 20 python tools/sanitize_small.py > gpurun_out/r06_san_$tool.log 2>&1; echo $tool rc=$?; tail -4 gpurun_out/r06_san_$tool.log
done
