echo "--- c2 8 GN, smoother blocks"; timeout 300 python tools/prof.py --workload c2 --lin 1 --pcg 20000 --gn 8 2>&1 | grep "optimize" | cut -c1-300
echo "--- c2 8 GN, 3x3 blocks"; TSB_SMOOTHER=0 timeout 300 python tools/prof.py --workload c2 --lin 1 --pcg 20000 --gn 8 2>&1 | grep "optimize" | cut -c1-300
echo "--- phase cycles smoother"; TSB_PCG_TIMING=1 timeout 300 python tools/prof.py --workload c2 --lin 1 --pcg 20000 2>&1 | grep "tsb\|solve" | cut -c1-400
echo "--- phase cycles 3x3"; TSB_SMOOTHER=0 TSB_PCG_TIMING=1 timeout 300 python tools/prof.py --workload c2 --lin 1 --pcg 20000 2>&1 | grep "tsb\|solve" | cut -c1-400
timeout 900 python -m pytest tests -m gpu -x -q -k "resident or oracle or reproducible or degenerate" > gpurun_out/pytest_gpu_sub.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu_sub.log | cut -c1-300
timeout 600 python bench.py --no-c4 > gpurun_out/bench_c2_sb.json 2> gpurun_out/bench_c2_sb.err; echo "bench rc=$?"; python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_c2_sb.json'))
print({k:d[k] for k in ('value','ms_per_step','gpu_launches')}, d['e2e']['value'], d['roofline']['frac'], d['config'].get('pcg_iters_per_gn'), d.get('solve_breakdown_ms_per_gn'))
PY
