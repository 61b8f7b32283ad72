timeout 600 python -m pytest tests -m gpu -x -q -k "streamed_pcg" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_gpu.log
timeout 600 python tools/prof.py --workload c4 --lin 1 --pcg 30 2>&1 | tail -1
TSB_NO_STREAM=1 timeout 600 python tools/prof.py --workload c4 --lin 1 --pcg 30 2>&1 | tail -1
timeout 600 python tools/prof.py --workload c3 --lin 1 --pcg 30 2>&1 | tail -1
TSB_NO_STREAM=1 timeout 600 python tools/prof.py --workload c3 --lin 1 --pcg 30 2>&1 | tail -1
