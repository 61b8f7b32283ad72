# round 2: the bench line at N ranks (one C3 request edge-sharded over the ranks, parity_vs_n1, C4 / C2 side keys)
N=${1:-2}
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29761 bench.py --gpus $N --steps 2 --warmup 1 > gpurun_out/bench_c3_n$N.json 2> gpurun_out/bench_c3_n$N.err; echo "bench rc=$?"
tail -5 gpurun_out/bench_c3_n$N.err | cut -c1-400
python - <<PY
import json
d = json.loads([l for l in open('gpurun_out/bench_c3_n$N.json') if l.startswith('{')][-1])
print({k: d[k] for k in ('value', 'ms_per_step', 'gpu_launches')}, 'e2e', d['e2e']['value'], 'roofline frac', d['roofline']['frac'], 'us/it', d['roofline']['us_per_pcg_iteration'], 'its/GN', d['config']['pcg_iters_per_gn'])
print('parity', d.get('parity_vs_n1')); print('lin', d['roofline_linearize']); print('c4', d.get('c4')); print('c2', d.get('c2_replicas')); print(d['solve_breakdown_ms_per_gn'])
PY
