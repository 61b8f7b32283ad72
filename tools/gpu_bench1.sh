# round 2: the bench line at N=1 (C3 headline, C4 / C2 side keys, CPU baseline)
timeout 1500 python bench.py --steps 2 --warmup 1 > gpurun_out/bench_c3_n1.json 2> gpurun_out/bench_c3_n1.err; echo "bench rc=$?"
tail -3 gpurun_out/bench_c3_n1.err | cut -c1-400
python - <<'PY'
import json
d = json.load(open('gpurun_out/bench_c3_n1.json'))
print({k: d[k] for k in ('value', 'ms_per_step', 'gpu_launches')}, 'e2e', d['e2e']['value'], 'roofline frac', d['roofline']['frac'], 'its/GN', d['config']['pcg_iters_per_gn'])
print('lin', d['roofline_linearize']); print('c4', d.get('c4')); print('c2', d.get('c2_replicas')); print('cpu', d.get('cpu_baseline')); print(d['solve_breakdown_ms_per_gn'])
PY
timeout 600 python bench.py --impl reference --steps 1 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; cut -c1-600 gpurun_out/bench_ref.json
