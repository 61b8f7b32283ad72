timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
timeout 600 python tools/prof.py --workload c3 --lin 1 --pcg 30 2>&1 | tail -1
TSB_NO_STREAM=1 timeout 600 python tools/prof.py --workload c3 --lin 1 --pcg 30 2>&1 | tail -1
timeout 600 python tools/prof.py --workload c5 --lin 1 --pcg 30 2>&1 | tail -1
TSB_NO_STREAM=1 timeout 600 python tools/prof.py --workload c5 --lin 1 --pcg 30 2>&1 | tail -1
timeout 900 python bench.py > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench rc=$?"; python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_c2.json'))
print({k:d[k] for k in ('value','ms_per_step')}, d['e2e']['value'])
print(d.get('roofline_linearize_c4')); print(d.get('roofline_pcg_c4'))
PY
