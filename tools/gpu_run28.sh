timeout 600 python -m pytest tests -m gpu -x -q -k "two_ranks" > gpurun_out/pytest_gpu_n2.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu_n2.log
echo "--- mgpu_check"; timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/mgpu_check.py 2>&1 | grep -v "^W\|warn\|Warn\|\*\*\*\|OMP_NUM" | tail -6 | tee gpurun_out/mgpu_check_n2.txt
echo "--- bench n2"; timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 > gpurun_out/bench_c2_n2.json 2> gpurun_out/bench_c2_n2.err; echo "bench rc=$?"; python - <<'PY'
import json
for ln in open('gpurun_out/bench_c2_n2.json'):
    if ln.startswith('{'):
        d=json.loads(ln)
        for k in ('value','e2e','single_request','roofline_linearize_c4'): print(k, d.get(k))
PY
