timeout 900 python -m pytest tests -m gpu -x -q -k "large_coarse or streamed or resident" > gpurun_out/pytest_gpu_sub.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu_sub.log | cut -c1-300
echo "--- c3 2 GN"; TSB_PCG_TIMING=1 timeout 600 python tools/prof.py --workload c3 --lin 1 --pcg 20000 --gn 2 2>&1 | grep "optimize\|streamed PCG" | cut -c1-400 | tail -2
echo "--- c4 1 GN"; TSB_PCG_TIMING=1 timeout 900 python tools/prof.py --workload c4 --lin 1 --pcg 20000 --gn 1 2>&1 | grep "optimize\|streamed PCG" | cut -c1-400 | tail -2
