timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -8 gpurun_out/pytest_gpu.log | cut -c1-300
echo "--- c4 phased 1 solve capped"; TSB_PCG_PHASED=1 timeout 600 python tools/prof.py --workload c4 --lin 1 --pcg 64 2>&1 | tail -1 | cut -c1-200
echo "--- c3 phased 1 solve capped"; TSB_PCG_PHASED=1 timeout 600 python tools/prof.py --workload c3 --lin 1 --pcg 64 2>&1 | tail -1 | cut -c1-200
echo "--- c3 persistent (no stream) 2 GN"; TSB_NO_STREAM=1 timeout 600 python tools/prof.py --workload c3 --lin 1 --pcg 20000 --gn 2 2>&1 | grep optimize | cut -c1-200
