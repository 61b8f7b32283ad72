timeout 900 python bench.py > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench rc=$?"; tail -2 gpurun_out/bench_c2.err; python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_c2.json'))
print({k:d[k] for k in ('value','ms_per_step')}, d['e2e']['value'])
print(d.get('roofline_linearize_c4')); print(d.get('roofline_pcg_c4'))
PY
TSB_PCG_TIMING=1 timeout 600 python tools/prof.py --workload c4 --lin 1 --pcg 30 2>&1 | grep "tsb\|solve" | tail -4
