# what the driver runs at round end, in its order: smoke, the GPU tests, the reference arm, the bench (N = 1)
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
( time timeout 900 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/final_ref_n1.json 2> gpurun_out/final_ref_n1.err ) 2>&1 | grep real
( time timeout 1500 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/final_bench_n1.json 2> gpurun_out/final_bench_n1.err ) 2>&1 | grep real
python - <<PY
import json
for f in ('gpurun_out/final_ref_n1.json', 'gpurun_out/final_bench_n1.json'):
    d = json.loads([l for l in open(f) if l.startswith('{')][-1])
    print(f, {k: d.get(k) for k in ('impl', 'value', 'steps', 'warmup', 'ms_per_step', 'gpu_launches')}, 'e2e', d['e2e']['value'], d.get('clocks'))
PY
