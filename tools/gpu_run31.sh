timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu.log
echo "--- GJ"; TSB_PCG_TIMING=1 timeout 300 python tools/prof.py --workload c2 --lin 1 --pcg 4000 --gn 3 2>&1 | grep "coarse setup" | tail -3
timeout 300 python tools/prof.py --workload c2 --lin 1 --pcg 4000 --gn 50 2>&1 | tail -1 | cut -c1-260
echo "--- cusolver"; TSB_COARSE_CUSOLVER=1 TSB_PCG_TIMING=1 timeout 300 python tools/prof.py --workload c2 --lin 1 --pcg 4000 --gn 3 2>&1 | grep "coarse setup" | tail -2
TSB_COARSE_CUSOLVER=1 timeout 300 python tools/prof.py --workload c2 --lin 1 --pcg 4000 --gn 50 2>&1 | tail -1 | cut -c1-260
