timeout 900 python bench.py > gpurun_out/bench_c2.json 2> gpurun_out/bench_c2.err; echo "bench rc=$?"; python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_c2.json'))
print({k:d[k] for k in ('value','ms_per_step','gpu_launches')}, d['e2e']['value'], d['roofline']['frac'])
print(d.get('roofline_linearize_c4')['frac'], d.get('roofline_pcg_c4')['frac'], d.get('roofline_pcg_c4')['us_per_pcg_iteration'])
PY
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv --log-file gpurun_out/launches_c2.csv python bench.py --steps 1 --warmup 1 --no-c4 --lin-reps 3 > gpurun_out/ncu_bench.log 2>&1; echo "ncu launches rc=$?"
echo "--- c5 3 GN"; timeout 900 python tools/prof.py --workload c5 --lin 1 --pcg 20000 --gn 3 2>&1 | grep "optimize" | cut -c1-300
timeout 300 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cut -c1-400 gpurun_out/bench_ref.json
