set -x
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
TSB_PCG_TIMING=1 timeout 600 python tools/prof.py --workload c2 --lin 2 --pcg 4000 --gn 5 2>&1 | tee gpurun_out/prof_c2_coarse.log
TSB_COARSE=0 timeout 600 python tools/prof.py --workload c2 --lin 2 --pcg 4000 --gn 5 2>&1 | tee gpurun_out/prof_c2_nocoarse.log
timeout 600 python tools/prof.py --workload c3 --lin 2 --pcg 20000 --gn 3 2>&1 | tee gpurun_out/prof_c3_coarse.log
timeout 900 python bench.py --steps 3 --warmup 3 --no-c4 > gpurun_out/bench_c2.log 2> gpurun_out/bench_c2.err; echo "bench rc=$?"; cat gpurun_out/bench_c2.log; tail -3 gpurun_out/bench_c2.err
