timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log | cut -c1-300
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench rc=$?"; python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_c2.json'))
print({k:d[k] for k in ('value','ms_per_step','gpu_launches')}, d['e2e']['value'], d['roofline']['frac'], d['config']['pcg_iters_per_gn'])
print(d.get('roofline_linearize_c4')['frac'], d.get('roofline_pcg_c4')['frac'], d.get('roofline_pcg_c4')['us_per_pcg_iteration'])
PY
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file gpurun_out/launches_c2.csv python bench.py --steps 1 --warmup 1 --no-c4 --lin-reps 3 > gpurun_out/ncu_bench.log 2>&1; echo "ncu launches rc=$?"
