timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
timeout 600 python tools/prof.py --workload c4 --lin 6 --pcg 0 2>&1 | tee gpurun_out/prof_c4_stream.log
TSB_NO_STREAM=1 timeout 600 python tools/prof.py --workload c4 --lin 4 --pcg 0 2>&1 | tail -2
timeout 600 python tools/prof.py --workload c3 --lin 6 --pcg 0 2>&1 | tail -3 | tee gpurun_out/prof_c3_stream.log
timeout 600 python tools/prof.py --workload c2 --lin 1 --pcg 4000 --gn 50 2>&1 | tail -1 | cut -c1-300
