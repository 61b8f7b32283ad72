echo "--- c2 x0.3 4 GN, 3x3 blocks"; timeout 300 python tools/prof.py --workload c2 --scale 0.3 --lin 1 --pcg 20000 --gn 4 2>&1 | grep "optimize" | cut -c1-300
echo "--- c2 x0.3 4 GN, smoother blocks"; TSB_SMOOTHER=1 timeout 300 python tools/prof.py --workload c2 --scale 0.3 --lin 1 --pcg 20000 --gn 4 2>&1 | grep "optimize" | cut -c1-300
timeout 900 python -m pytest tests -m gpu -x -q -k "large_coarse or resident" > gpurun_out/pytest_gpu_sub.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu_sub.log | cut -c1-300
echo "--- c3 3 GN"; TSB_PCG_TIMING=1 timeout 600 python tools/prof.py --workload c3 --lin 1 --pcg 20000 --gn 3 2>&1 | grep "optimize\|coarse setup" | cut -c1-300 | tail -5
echo "--- c3 3 GN cusolver"; TSB_COARSE_CUSOLVER=1 TSB_PCG_TIMING=1 timeout 600 python tools/prof.py --workload c3 --lin 1 --pcg 20000 --gn 3 2>&1 | grep "optimize\|coarse setup" | cut -c1-300 | tail -3
echo "--- c4 2 GN"; TSB_PCG_TIMING=1 timeout 900 python tools/prof.py --workload c4 --lin 1 --pcg 20000 --gn 2 2>&1 | grep "optimize\|coarse setup\|streamed PCG" | cut -c1-400 | tail -6
echo "--- c5 2 GN"; timeout 900 python tools/prof.py --workload c5 --lin 1 --pcg 20000 --gn 2 2>&1 | grep "optimize" | cut -c1-300
