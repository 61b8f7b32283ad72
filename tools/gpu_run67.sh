echo "--- c2 50 GN adaptive"; timeout 300 python tools/prof.py --workload c2 --lin 1 --pcg 20000 --gn 50 2>&1 | grep "optimize" | cut -c1-330
timeout 900 python -m pytest tests -m gpu -x -q -k "resident or full_size or reproducible or oracle" 2>&1 | tail -2
timeout 900 python bench.py --no-c4 > gpurun_out/bench_c2_b.json 2> gpurun_out/bench_c2.err; echo "bench rc=$?"; python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_c2_b.json'))
print({k:d[k] for k in ('value','ms_per_step','gpu_launches')}, d['e2e']['value'], d['config']['pcg_iters_per_gn'])
PY
