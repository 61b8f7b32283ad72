echo "--- c2 50 GN adaptive"; timeout 300 python tools/prof.py --workload c2 --lin 1 --pcg 20000 --gn 50 2>&1 | grep "optimize" | cut -c1-900
echo "--- c2 50 GN 3x3"; TSB_SMOOTHER=0 timeout 300 python tools/prof.py --workload c2 --lin 1 --pcg 20000 --gn 50 2>&1 | grep "optimize" | cut -c1-900
echo "--- c2 50 GN sb"; TSB_SMOOTHER=1 timeout 300 python tools/prof.py --workload c2 --lin 1 --pcg 20000 --gn 50 2>&1 | grep "optimize" | cut -c1-900
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log | cut -c1-300
timeout 900 python bench.py --no-c4 > gpurun_out/bench_c2_adaptive.json 2> gpurun_out/bench_c2.err; echo "bench rc=$?"; python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_c2_adaptive.json'))
print({k:d[k] for k in ('value','ms_per_step','gpu_launches')}, d['e2e']['value'], d['roofline']['frac'], d['config']['pcg_iters_per_gn'], d['solve_breakdown_ms_per_gn'])
PY
