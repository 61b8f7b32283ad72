timeout 900 python -m pytest tests -m gpu -x -q -k "phased or streamed or more_edges" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_gpu.log | cut -c1-200
echo "--- c4 phased, streamed passes"; TSB_PCG_PHASED=1 timeout 600 python tools/prof.py --workload c4 --lin 1 --pcg 32 2>&1 | tail -1
echo "--- c4 phased, unit passes"; TSB_PCG_PHASED=1 TSB_NO_STREAM=1 timeout 600 python tools/prof.py --workload c4 --lin 1 --pcg 32 2>&1 | tail -1
