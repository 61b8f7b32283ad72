timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 600 python tools/prof.py --workload c4 --lin 5 --pcg 0 2>&1 | tail -2
TSB_PCG_TIMING=1 timeout 600 python tools/prof.py --workload c4 --lin 1 --pcg 30 2>&1 | grep "persistent\|solve" | tail -3
timeout 600 python tools/prof.py --workload c3 --lin 3 --pcg 30 2>&1 | tail -2
