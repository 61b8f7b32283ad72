echo "--- bench n2"; timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 > gpurun_out/bench_c2_n2.json 2> gpurun_out/bench_c2_n2.err; echo "bench rc=$?"; grep -v "^\*\|OMP_NUM" gpurun_out/bench_c2_n2.err | tail -5; python - <<'PY'
import json
for ln in open('gpurun_out/bench_c2_n2.json'):
    if ln.startswith('{'):
        d=json.loads(ln)
        for k in ('value','e2e','single_request','roofline_linearize_c4','roofline_pcg_c4'): print(k, d.get(k))
PY
