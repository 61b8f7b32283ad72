timeout 600 python -m pytest tests -m gpu -x -q -k "streaming or structure or golden or sparse" > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
echo "--- c4"; timeout 600 python tools/prof.py --workload c4 --lin 6 --pcg 0 2>&1 | tail -2 | tee gpurun_out/prof_c4_planes.log
echo "--- c3"; timeout 600 python tools/prof.py --workload c3 --lin 6 --pcg 0 2>&1 | tail -2
echo "--- c5"; timeout 600 python tools/prof.py --workload c5 --lin 6 --pcg 0 2>&1 | tail -2
