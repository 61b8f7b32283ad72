timeout 600 python -m pytest tests -m gpu -x -q -k "streamed_pcg" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
TSB_PCG_TIMING=1 timeout 600 python tools/prof.py --workload c4 --lin 1 --pcg 30 2>&1 | grep "streamed\|solve" | tail -3
