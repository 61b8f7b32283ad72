# compute-sanitizer over the streamed kernels (TMA ring: sources aligned down / sizes rounded up to 16 B) on a graph whose
# LM edge count is not a multiple of 4 (19 edges per pose, 1301 poses), single GPU; memcheck then racecheck
export TSB_NO_RESIDENT=1 TSB_STREAM_MIN_UNITS=1
for tool in memcheck racecheck; do
  echo "=== compute-sanitizer --tool $tool"
  timeout 1200 compute-sanitizer --tool $tool --error-exitcode 9 python tools/prof.py --workload c4 --scale 0.001301 --lin 2 --pcg 25 --gn 1 > gpurun_out/sanitize_$tool.log 2>&1; echo "rc=$?"
  grep -c "Invalid\|Race\|hazard" gpurun_out/sanitize_$tool.log; grep "ERROR SUMMARY\|RACECHECK SUMMARY\|linearize\|solve\|optimize" gpurun_out/sanitize_$tool.log | tail -5 | cut -c1-200
done
