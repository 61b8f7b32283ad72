echo "--- bench n2"; timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 > gpurun_out/bench_c2_n2.json 2> gpurun_out/bench_c2_n2.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_c2_n2.err; python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_c2_n2.json'))
for k in ('value','scaling','ms_per_step','e2e','single_request','roofline_linearize_c4'): print(k, d.get(k))
print(d['config']['parallelism'])
PY
echo "--- ref n2"; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus 2 --steps 1 --warmup 1 2>/dev/null | tail -1 | cut -c1-400
