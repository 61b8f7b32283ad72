# bench.py at N ranks (N = $1), the way the driver launches it
N=$1
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29540 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/bench_c2_n$N.json 2> gpurun_out/bench_c2_n$N.err; echo "bench n$N rc=$?"; tail -2 gpurun_out/bench_c2_n$N.err | cut -c1-300
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/bench_c2_n$N.json') if l.startswith('{')][-1])
print({k:d[k] for k in ('value','ms_per_step','n_gpus','scaling')}, d['e2e']['value'])
print(d.get('single_request'))
for k in ('roofline_linearize_c4','roofline_pcg_c4'):
    r=d.get(k) or {}; print(k, {q:r.get(q) for q in ('frac','ms','us_per_pcg_iteration','edges_linearized_per_s','error')})
PY
