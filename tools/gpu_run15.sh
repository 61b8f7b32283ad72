timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -7 gpurun_out/pytest_gpu.log
timeout 300 python tools/prof.py --workload c2 --lin 1 --pcg 4000 --gn 50 2>&1 | tail -1 | cut -c1-400
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_c2.log 2> gpurun_out/bench_c2.err; echo "bench rc=$?"; cat gpurun_out/bench_c2.log | cut -c1-1200; tail -3 gpurun_out/bench_c2.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.log 2>&1; cat gpurun_out/bench_ref.log | cut -c1-600
