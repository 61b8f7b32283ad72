set -x
nvidia-smi -L
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" 
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_c2.log 2> gpurun_out/bench_c2.err; echo "bench rc=$?"
timeout 600 python tools/prof.py --workload c4 --lin 5 --pcg 20 > gpurun_out/prof_c4.log 2>&1; echo "prof rc=$?"
timeout 600 python tools/prof.py --workload c3 --lin 5 --pcg 50 --gn 5 > gpurun_out/prof_c3.log 2>&1; echo "prof3 rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_c2.csv python bench.py --steps 1 --warmup 1 --no-c4 --lin-reps 3 > gpurun_out/ncu_bench.log 2>&1; echo "ncu rc=$?"
tail -5 gpurun_out/pytest_gpu.log; cat gpurun_out/bench_c2.log; cat gpurun_out/prof_c4.log gpurun_out/prof_c3.log
