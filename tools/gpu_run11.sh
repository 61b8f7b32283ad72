timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -7 gpurun_out/pytest_gpu.log
TSB_PCG_TIMING=1 timeout 600 python tools/prof.py --workload c2 --lin 1 --pcg 4000 --gn 2 2>&1 | grep coarse | tail -2
timeout 600 python tools/prof.py --workload c2 --lin 1 --pcg 4000 --gn 50 2>&1 | tail -1 | cut -c1-700
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_c2.log 2> gpurun_out/bench_c2.err; echo "bench rc=$?"; cat gpurun_out/bench_c2.log; tail -3 gpurun_out/bench_c2.err
