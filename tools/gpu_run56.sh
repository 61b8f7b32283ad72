timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu.log | cut -c1-300
echo "--- c3 3 GN"; TSB_PCG_TIMING=1 timeout 600 python tools/prof.py --workload c3 --lin 1 --pcg 20000 --gn 3 2>&1 | grep "optimize\|coarse setup\|streamed PCG" | cut -c1-400 | tail -3
echo "--- c4 2 GN"; TSB_PCG_TIMING=1 timeout 900 python tools/prof.py --workload c4 --lin 1 --pcg 20000 --gn 2 2>&1 | grep "optimize\|coarse setup\|streamed PCG" | cut -c1-400 | tail -3
echo "--- c4 phased 1 solve capped"; TSB_PCG_PHASED=1 timeout 600 python tools/prof.py --workload c4 --lin 1 --pcg 64 2>&1 | tail -1 | cut -c1-200
timeout 900 python bench.py > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench rc=$?"; python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_c2.json'))
print({k:d[k] for k in ('value','ms_per_step','gpu_launches')}, d['e2e']['value'], d['roofline']['frac'])
print(d.get('roofline_linearize_c4')['frac'], d.get('roofline_pcg_c4')['frac'], d.get('roofline_pcg_c4')['us_per_pcg_iteration'])
PY
