echo "--- c3 phased, only A streamed"; TSB_PS_ONLY=a TSB_PCG_PHASED=1 CUDA_LAUNCH_BLOCKING=1 timeout 600 python tools/prof.py --workload c3 --lin 1 --pcg 32 2>&1 | tail -1 | cut -c1-200
echo "--- c3 phased, only B streamed"; TSB_PS_ONLY=b TSB_PCG_PHASED=1 CUDA_LAUNCH_BLOCKING=1 timeout 600 python tools/prof.py --workload c3 --lin 1 --pcg 32 2>&1 | tail -1 | cut -c1-200
