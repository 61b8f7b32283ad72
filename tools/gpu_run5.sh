timeout 600 python tools/prof.py --workload c2 --lin 1 --pcg 4000 --gn 50 2>&1 | tail -1 | tee gpurun_out/prof_c2_lag8.log
TSB_COARSE_MAX_AGE=1 timeout 600 python tools/prof.py --workload c2 --lin 1 --pcg 4000 --gn 50 2>&1 | tail -1 | tee gpurun_out/prof_c2_lag1.log
TSB_COARSE_MAX_AGE=50 timeout 600 python tools/prof.py --workload c2 --lin 1 --pcg 4000 --gn 50 2>&1 | tail -1 | tee gpurun_out/prof_c2_lag50.log
timeout 600 python tools/prof.py --workload c3 --lin 1 --pcg 20000 --gn 10 2>&1 | tail -1 | tee gpurun_out/prof_c3_lag8.log
timeout 900 python bench.py --steps 3 --warmup 3 --no-c4 > gpurun_out/bench_c2.log 2> gpurun_out/bench_c2.err; echo "bench rc=$?"; cat gpurun_out/bench_c2.log | cut -c1-900
